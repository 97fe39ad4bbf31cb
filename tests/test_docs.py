"""The documents quote evidence by file name: every profiles/ / tools/ / tests/ path they mention must exist."""
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_paths_quoted_in_the_documents_exist():
    missing = []
    for doc in ("DESIGN.md", "README.md", "INTEGRATION.md", os.path.join("profiles", "README.md")):
        text = open(os.path.join(ROOT, doc)).read()
        for m in re.finditer(r"`((?:profiles|tools|tests|integration|oracle|hevc_b200|include)/[A-Za-z0-9_./{},*-]+)`", text):
            path = m.group(1).rstrip(".,")
            if "{" in path:                                    # r02_final_scale{2,4,8}_bench_4k.json
                head, rest = path.split("{", 1)
                alts, tail = rest.split("}", 1)
                paths = [head + a + tail for a in alts.split(",")]
            else:
                paths = [path]
            for p in paths:
                if "*" in p:
                    ok = bool(glob.glob(os.path.join(ROOT, p)))
                else:
                    ok = os.path.exists(os.path.join(ROOT, p))
                # build outputs are not in the tree before build()
                if not ok and not p.endswith(".so") and "_ref/" not in p and "_build/" not in p and "_ubench" not in p:
                    missing.append(f"{doc}: {p}")
    assert not missing, "\n".join(missing)
