#!/bin/bash
# rebuild every native piece from anywhere
cd /root/repo && python -c "import __graft_entry__ as g; g.build()"
