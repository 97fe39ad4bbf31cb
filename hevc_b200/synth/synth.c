/*
 * synth.c — deterministic synthetic work lists (see ohsynth.h).
 * Test / bench input generator; mirrors what the reference control plane emits:
 * hls_coding_quadtree / hls_coding_unit / hls_prediction_unit / hls_transform_tree
 * (hevcdec.c:2711, :2550, :2116, :1549) without a bitstream.
 */
#include "ohsynth.h"

#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ PRNG */
typedef struct rng { uint64_t s[4]; } rng;
static uint64_t splitmix64(uint64_t *x)
{
    uint64_t z = (*x += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static void rng_seed(rng *r, uint64_t seed)
{
    for (int i = 0; i < 4; i++) r->s[i] = splitmix64(&seed);
}
static inline uint64_t rotl64(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
static uint64_t rnd(rng *r)                 /* xoshiro256** */
{
    uint64_t *s = r->s, res = rotl64(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl64(s[3], 45);
    return res;
}
static int rnd_range(rng *r, int lo, int hi) { return lo + (int)((rnd(r) >> 11) % (uint64_t)(hi - lo + 1)); }
static int rnd_pct(rng *r, int pct) { return (int)((rnd(r) >> 11) % 100) < pct; }
/* two-sided geometric-ish magnitude with scale ~sigma */
static int rnd_laplace(rng *r, int sigma)
{
    int a = rnd_range(r, 0, sigma), b = rnd_range(r, 0, sigma), c = rnd_range(r, 0, 3);
    int m = c == 0 ? a + b : (c == 1 ? a : a >> 2);
    return (rnd(r) >> 20 & 1) ? -m : m;
}

/* ------------------------------------------------------------------ state */
typedef struct gen {
    const ohsynth_params *p;
    ohsynth_frame *f;
    rng r;
    int hs, vs, cfi;
    int W, H, bd;
    int cap_pu, cap_tu, cap_iu;
    int64_t cap_coef;
    uint8_t *done;          /* per 4x4 luma cell: reconstructed already (intra availability) */
    int cell_w, cell_h;
    int intra;              /* the CU being generated is intra */
    int slice;              /* current slice index */
    int qp;
    uint16_t scan_rank[4][32 * 32]; /* [log2-2][y*n+x] = position in coding order */
} gen;

static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int iclip(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* Coding order of an n x n TB: 4x4 coefficient groups in up-right diagonal order,
 * and the same diagonal order inside each group (the inter default, SCAN_DIAG). */
static void build_scan(gen *g)
{
    int d4[16], k = 0;
    for (int d = 0; d < 7; d++)
        for (int y = imin(d, 3); y >= 0 && d - y <= 3; y--)
            d4[k++] = y * 4 + (d - y);
    for (int l = 0; l < 4; l++) {
        int n = 4 << l, ncg = n >> 2, rank = 0;
        for (int d = 0; d < 2 * ncg - 1; d++)
            for (int cy = imin(d, ncg - 1); cy >= 0 && d - cy <= ncg - 1; cy--) {
                int cx = d - cy;
                for (int i = 0; i < 16; i++) {
                    int x = cx * 4 + (d4[i] & 3), y = cy * 4 + (d4[i] >> 2);
                    g->scan_rank[l][y * n + x] = (uint16_t)rank++;
                }
            }
    }
}

static void push_pu(gen *g, const ohcu_pu *pu)
{
    ohsynth_frame *f = g->f;
    if (f->n_pu == g->cap_pu) {
        g->cap_pu = g->cap_pu ? g->cap_pu * 2 : 4096;
        f->pus = realloc(f->pus, sizeof(ohcu_pu) * g->cap_pu);
    }
    f->pus[f->n_pu++] = *pu;
}
static int16_t *push_tu(gen *g, ohcu_tu *tu)
{
    ohsynth_frame *f = g->f;
    int n2 = 1 << (2 * tu->log2_size);
    if (f->n_tu == g->cap_tu) {
        g->cap_tu = g->cap_tu ? g->cap_tu * 2 : 4096;
        f->tus = realloc(f->tus, sizeof(ohcu_tu) * g->cap_tu);
    }
    while (f->ncoeff + n2 > g->cap_coef) {
        g->cap_coef = g->cap_coef ? g->cap_coef * 2 : 1 << 20;
        f->coeffs = realloc(f->coeffs, sizeof(int16_t) * g->cap_coef);
    }
    tu->coeff_off = (uint32_t)f->ncoeff;
    f->tus[f->n_tu++] = *tu;
    memset(f->coeffs + f->ncoeff, 0, sizeof(int16_t) * n2);
    f->ncoeff += n2;
    return f->coeffs + tu->coeff_off;
}

/* boundary strength on the left and top edges of a luma block (8x8 grid only) */
static void paint_bs(gen *g, int x, int y, int w, int h)
{
    ohsynth_frame *f = g->f;
    const int *pc = g->p->bs_pct;
    if ((x & 7) == 0)
        for (int yy = y; yy < y + h; yy += 4) {
            int v = rnd_range(&g->r, 0, 99);
            f->vbs[(x + yy * f->bs_width) >> 2] = v < pc[0] ? 1 : (v < pc[0] + pc[1] ? 2 : 0);
        }
    if ((y & 7) == 0)
        for (int xx = x; xx < x + w; xx += 4) {
            int v = rnd_range(&g->r, 0, 99);
            f->hbs[(xx + y * f->bs_width) >> 2] = v < pc[0] ? 1 : (v < pc[0] + pc[1] ? 2 : 0);
        }
}

/* ------------------------------------------------------------------ PUs */
static void emit_pu(gen *g, int x, int y, int w, int h)
{
    const ohsynth_params *p = g->p;
    ohcu_pu pu;
    memset(&pu, 0, sizeof(pu));
    pu.x0 = (uint16_t)x; pu.y0 = (uint16_t)y; pu.w = (uint8_t)w; pu.h = (uint8_t)h;
    pu.slice = (uint8_t)g->slice;
    /* 8x4 / 4x8 are never bi-predicted (hevc_mvs.c:549-550) */
    if (w + h != 12 && rnd_pct(&g->r, p->bi_pct))
        pu.pred_flag = OHCU_PF_BI;
    else
        pu.pred_flag = rnd_pct(&g->r, 50) ? OHCU_PF_L0 : OHCU_PF_L1;
    for (int l = 0; l < 2; l++) {
        int far = rnd_pct(&g->r, p->far_mv_pct);
        int span = far ? 2048 * 4 : 64 * 4;
        pu.mv[l][0] = (int16_t)rnd_range(&g->r, -span, span);
        pu.mv[l][1] = (int16_t)rnd_range(&g->r, -span, span);
        /* a share of exact full-pel / one-axis vectors so every [!!my][!!mx] occurs */
        int k = rnd_range(&g->r, 0, 9);
        if (k == 0) { pu.mv[l][0] &= ~3; pu.mv[l][1] &= ~3; }
        else if (k == 1) pu.mv[l][0] &= ~3;
        else if (k == 2) pu.mv[l][1] &= ~3;
        pu.ref_idx[l] = (uint8_t)rnd_range(&g->r, 0, p->n_refs - 1);
    }
    push_pu(g, &pu);
    paint_bs(g, x, y, w, h);
}

/* ------------------------------------------------------------------ TUs */
/* fill one coded TB; returns the index of the record */
static int emit_tb(gen *g, int c_idx, int x, int y, int log2, int bypass, int cross_from)
{
    const ohsynth_params *p = g->p;
    int n = 1 << log2, l = log2 - 2;
    ohcu_tu tu;
    int16_t *c;
    memset(&tu, 0, sizeof(tu));
    tu.x0 = (uint16_t)x; tu.y0 = (uint16_t)y; tu.c_idx = (uint8_t)c_idx; tu.log2_size = (uint8_t)log2;
    if (g->intra) tu.flags |= OHCU_TUF_INTRA;
    if (cross_from >= 0) {
        static const int8_t scale[8] = { 1, 2, 4, 8, -1, -2, -4, -8 };
        tu.flags |= OHCU_TUF_CROSS;
        tu.res_scale_val = scale[rnd_range(&g->r, 0, 7)];
        tu.cross_tu = (uint32_t)cross_from;
    }
    if (bypass) {
        tu.mode = OHCU_TU_BYPASS;
        if (p->rext && rnd_pct(&g->r, 30))
            tu.flags |= OHCU_TUF_RDPCM | (rnd_pct(&g->r, 50) ? OHCU_TUF_RDPCM_VERT : 0);
        c = push_tu(g, &tu);
        for (int i = 0; i < n * n; i++) c[i] = (int16_t)rnd_laplace(&g->r, 24);
    } else if (p->rext && rnd_pct(&g->r, 8)) {
        tu.mode = OHCU_TU_SKIP;
        if (log2 == 2 && rnd_pct(&g->r, 50)) tu.flags |= OHCU_TUF_ROTATE;
        if (rnd_pct(&g->r, 30))
            tu.flags |= OHCU_TUF_RDPCM | (rnd_pct(&g->r, 50) ? OHCU_TUF_RDPCM_VERT : 0);
        c = push_tu(g, &tu);
        for (int i = 0; i < n * n; i++)
            c[i] = (int16_t)(rnd_pct(&g->r, 2) ? rnd_range(&g->r, -32768, 32767) : rnd_laplace(&g->r, 300));
    } else if (p->dense_coeffs) {
        /* hazard 1: dense block, arbitrary col_limit (exercises the zero-skip masks) */
        tu.mode = (c_idx == 0 && log2 == 2 && (g->intra || rnd_pct(&g->r, p->dst_pct))) ? OHCU_TU_DST4 : OHCU_TU_IDCT;
        tu.col_limit = (int8_t)rnd_range(&g->r, 1, n + 8);
        c = push_tu(g, &tu);
        for (int i = 0; i < n * n; i++)
            c[i] = (int16_t)(rnd_pct(&g->r, 3) ? rnd_range(&g->r, -32768, 32767) : rnd_laplace(&g->r, 400));
    } else {
        /* 4x4 luma blocks of intra CUs always take the DST (hevc_cabac.c:1790); dst_pct spreads it over inter CUs too */
        int lx, ly, last_rank, dst4 = c_idx == 0 && log2 == 2 && (g->intra || rnd_pct(&g->r, p->dst_pct));
        if (rnd_pct(&g->r, p->dc_pct)) {
            lx = ly = 0;
        } else {
            /* bias the last significant position towards low frequencies */
            int m = rnd_pct(&g->r, 60) ? imin(n, 8) : n;
            lx = rnd_range(&g->r, 0, m - 1); ly = rnd_range(&g->r, 0, m - 1);
        }
        last_rank = g->scan_rank[l][ly * n + lx];
        if (dst4) {
            tu.mode = OHCU_TU_DST4;
        } else if (lx == 0 && ly == 0) {
            tu.mode = OHCU_TU_DC;                              /* hevc_cabac.c:1794-1796 */
        } else {
            int max_xy = lx > ly ? lx : ly, lim = lx + ly + 4;  /* hevc_cabac.c:1797-1805 */
            if (max_xy < 4) lim = imin(4, lim);
            else if (max_xy < 8) lim = imin(8, lim);
            else if (max_xy < 12) lim = imin(24, lim);
            tu.mode = OHCU_TU_IDCT;
            tu.col_limit = (int8_t)lim;
        }
        c = push_tu(g, &tu);
        for (int i = 0; i < n * n; i++) {
            if (g->scan_rank[l][i] > last_rank) continue;
            if (g->scan_rank[l][i] != last_rank && !rnd_pct(&g->r, 45)) continue;
            c[i] = (int16_t)(rnd_pct(&g->r, 1) ? rnd_range(&g->r, -32768, 32767) : rnd_laplace(&g->r, 200));
        }
        if (c[ly * n + lx] == 0) c[ly * n + lx] = 1;
    }
    return g->f->n_tu - 1;
}


/* ------------------------------------------------------------------ intra blocks */
static void mark_done(gen *g, int x, int y, int w, int h)
{
    for (int yy = y >> 2; yy < (y + h) >> 2 && yy < g->cell_h; yy++)
        for (int xx = x >> 2; xx < (x + w) >> 2 && xx < g->cell_w; xx++)
            g->done[yy * g->cell_w + xx] = 1;
}
/* the 4x4 luma cell at (xn, yn) was reconstructed before the block at (x, y) and may be read by it: inside the
 * picture, earlier in decoding order, same slice and same tile (what lc->na.cand_* and the z-scan test of
 * hevcpred_template.c:103-107 amount to) */
static int avail(const gen *g, int x, int y, int xn, int yn)
{
    const ohsynth_frame *f = g->f;
    int l2 = g->p->sp.log2_ctb_size, a, b;
    if (xn < 0 || yn < 0 || xn >= g->W || yn >= g->H) return 0;
    if (!g->done[(yn >> 2) * g->cell_w + (xn >> 2)]) return 0;
    a = (y >> l2) * f->ctb_width + (x >> l2); b = (yn >> l2) * f->ctb_width + (xn >> l2);
    return f->tab_slice_address[a] == f->tab_slice_address[b] && f->tile_id[a] == f->tile_id[b];
}
/* one predicted block of plane c_idx at plane position (x, y); has_res: a coded TB follows */
static void emit_iu(gen *g, int c_idx, int x, int y, int log2, int mode, int has_res)
{
    ohsynth_frame *f = g->f;
    ohcu_iu iu;
    int hs = c_idx ? g->hs : 0, vs = c_idx ? g->vs : 0;
    int xl = x << hs, yl = y << vs, wl = (1 << log2) << hs, hl = (1 << log2) << vs, cand = 0;
    if (avail(g, xl, yl, xl - 1, yl + hl)) cand |= OHCU_CAND_BOTTOM_LEFT;
    if (avail(g, xl, yl, xl - 1, yl))      cand |= OHCU_CAND_LEFT;
    if (avail(g, xl, yl, xl - 1, yl - 1))  cand |= OHCU_CAND_UP_LEFT;
    if (avail(g, xl, yl, xl, yl - 1))      cand |= OHCU_CAND_UP;
    if (avail(g, xl, yl, xl + wl, yl - 1)) cand |= OHCU_CAND_UP_RIGHT;
    if (g->p->intra_drop_pct && rnd_pct(&g->r, g->p->intra_drop_pct)) cand &= rnd_range(&g->r, 0, 31);
    memset(&iu, 0, sizeof(iu));
    iu.x0 = (uint16_t)x; iu.y0 = (uint16_t)y; iu.c_idx = (uint8_t)c_idx; iu.log2_size = (uint8_t)log2;
    iu.mode = (uint8_t)mode; iu.cand = (uint8_t)(cand | (has_res ? OHCU_IU_CBF : 0));
    if (f->n_iu == g->cap_iu) {
        g->cap_iu = g->cap_iu ? g->cap_iu * 2 : 4096;
        f->ius = realloc(f->ius, sizeof(ohcu_iu) * g->cap_iu);
    }
    f->ius[f->n_iu++] = iu;
}
static int draw_mode(gen *g)
{
    int k = rnd_range(&g->r, 0, 9);        /* planar, DC and the two pure directions more often */
    return k == 0 ? 0 : k == 1 ? 1 : k == 2 ? 10 : k == 3 ? 26 : rnd_range(&g->r, 0, 34);
}
static void emit_tb_or_cross(gen *g, int c, int x, int y, int log2, int bypass, int cross, int coded);
/* chroma blocks of an intra TU, luma position (x, y), luma size 1 << log2 (hls_transform_unit, hevcdec.c:1380-1526) */
static void emit_chroma_intra(gen *g, int x, int y, int log2, int bypass, int luma_idx, int mode_c)
{
    int cfi = g->cfi;
    if (!cfi) return;
    for (int c = 1; c <= 2; c++) {
        if (cfi == 3) {
            int cross = (g->p->rext && luma_idx >= 0 && rnd_pct(&g->r, 25)) ? luma_idx : -1;
            int coded = rnd_pct(&g->r, g->p->cbf_pct);
            emit_iu(g, c, x, y, log2, mode_c, coded || cross >= 0);
            emit_tb_or_cross(g, c, x, y, log2, bypass, cross, coded);
        } else {
            int l2c = log2 - 1, parts = cfi == 2 ? 2 : 1;
            for (int i = 0; i < parts; i++) {
                int coded = rnd_pct(&g->r, g->p->cbf_pct);
                emit_iu(g, c, x >> 1, (y >> g->vs) + (i << l2c), l2c, mode_c, coded);
                if (coded) emit_tb(g, c, x >> 1, (y >> g->vs) + (i << l2c), l2c, bypass, -1);
            }
        }
    }
}

/* chroma TBs belonging to a luma block at (x,y) of size 1<<log2 (luma units) */
static void emit_chroma(gen *g, int x, int y, int log2, int bypass, int luma_idx)
{
    int cfi = g->cfi;
    if (!cfi) return;
    for (int c = 1; c <= 2; c++) {
        if (cfi == 3) {
            int cross = (g->p->rext && luma_idx >= 0 && rnd_pct(&g->r, 25)) ? luma_idx : -1;
            emit_tb_or_cross(g, c, x, y, log2, bypass, cross, rnd_pct(&g->r, g->p->cbf_pct));
        } else {
            int l2c = log2 - 1, parts = cfi == 2 ? 2 : 1;
            for (int i = 0; i < parts; i++)
                if (rnd_pct(&g->r, g->p->cbf_pct))
                    emit_tb(g, c, x >> 1, (y >> g->vs) + (i << l2c), l2c, bypass, -1);
        }
    }
}

static void emit_tb_or_cross(gen *g, int c, int x, int y, int log2, int bypass, int cross, int coded)
{
    if (coded)
        emit_tb(g, c, x, y, log2, bypass, cross);
    else if (cross >= 0) {
        /* uncoded chroma with cross-component prediction (hevcdec.c:1430-1445):
         * residual = (res_scale_val * coeffs_y) >> 3 on top of zero coefficients */
        int idx = emit_tb(g, c, x, y, log2, 1, cross);
        ohcu_tu *tu = &g->f->tus[idx];
        tu->flags &= ~(OHCU_TUF_RDPCM | OHCU_TUF_RDPCM_VERT);
        memset(g->f->coeffs + tu->coeff_off, 0, sizeof(int16_t) << (2 * log2));
    }
}

/* transform tree of an intra CU: every leaf is predicted (luma, then chroma), coded or not.
 * modes[4]: luma mode per quadrant of the CU (NxN) or four times the same; force: NxN splits at depth 0 */
static void gen_tt_intra(gen *g, int x, int y, int log2, int bypass, const int *modes, int mode_c, int force, int quad)
{
    int split = log2 > 5 || force || (log2 > 2 && rnd_pct(&g->r, 35));
    if (split) {
        int h = 1 << (log2 - 1);
        for (int i = 0; i < 4; i++)
            gen_tt_intra(g, x + (i & 1) * h, y + (i >> 1) * h, log2 - 1, bypass, modes, mode_c, 0, force ? i : quad);
        if (log2 == 3 && g->cfi && g->cfi != 3)
            emit_chroma_intra(g, x, y, 3, bypass, -1, mode_c);
        return;
    }
    paint_bs(g, x, y, 1 << log2, 1 << log2);
    {
        int luma_idx = -1, coded = rnd_pct(&g->r, g->p->cbf_pct);
        ohsynth_leaf *lf = &g->f->leaves[g->f->n_leaf++];
        emit_iu(g, 0, x, y, log2, modes[quad], coded);
        if (coded)
            luma_idx = emit_tb(g, 0, x, y, log2, bypass, -1);
        mark_done(g, x, y, 1 << log2, 1 << log2);
        lf->x0 = (uint16_t)x; lf->y0 = (uint16_t)y; lf->log2_size = (uint8_t)log2; lf->cbf_luma = luma_idx >= 0;
        if (log2 > 2 || g->cfi == 3)
            emit_chroma_intra(g, x, y, log2, bypass, luma_idx, mode_c);
    }
}

/* hls_transform_tree, hevcdec.c:1549 */
static void gen_tt(gen *g, int x, int y, int log2, int bypass)
{
    int split = log2 > 5 || (log2 > 2 && rnd_pct(&g->r, 35));
    if (split) {
        int h = 1 << (log2 - 1);
        for (int i = 0; i < 4; i++)
            gen_tt(g, x + (i & 1) * h, y + (i >> 1) * h, log2 - 1, bypass);
        /* four 4x4 luma TBs share the chroma TBs of their 8x8 parent when chroma
         * is horizontally subsampled (blk_idx == 3 branch, hevcdec.c:1480-1526) */
        if (log2 == 3 && g->cfi && g->cfi != 3)
            emit_chroma(g, x, y, 3, bypass, -1);
        return;
    }
    paint_bs(g, x, y, 1 << log2, 1 << log2);
    {
        int luma_idx = -1;
        ohsynth_leaf *lf = &g->f->leaves[g->f->n_leaf++];
        if (rnd_pct(&g->r, g->p->cbf_pct))
            luma_idx = emit_tb(g, 0, x, y, log2, bypass, -1);
        lf->x0 = (uint16_t)x; lf->y0 = (uint16_t)y; lf->log2_size = (uint8_t)log2; lf->cbf_luma = luma_idx >= 0;
        if (log2 > 2 || g->cfi == 3)
            emit_chroma(g, x, y, log2, bypass, luma_idx);
    }
}

/* hls_coding_unit (inter), hevcdec.c:2550 */
static void gen_cu(gen *g, int x, int y, int log2)
{
    const ohsynth_params *p = g->p;
    ohsynth_frame *f = g->f;
    int s = 1 << log2, q = s >> 2, hlf = s >> 1;
    int bypass = p->sp.pcm_or_bypass && rnd_pct(&g->r, p->bypass_pct);
    int mode;

    g->qp = iclip(g->qp + rnd_range(&g->r, -3, 3), 18, 45);
    for (int yy = y >> p->sp.log2_min_cb_size; yy < (y + s) >> p->sp.log2_min_cb_size; yy++)
        for (int xx = x >> p->sp.log2_min_cb_size; xx < (x + s) >> p->sp.log2_min_cb_size; xx++)
            f->qp_y_tab[yy * f->min_cb_width + xx] = (int8_t)g->qp;
    if (bypass)                        /* set_deblocking_bypass, hevcdec.c */
        for (int yy = y >> p->sp.log2_min_pu_size; yy < (y + s) >> p->sp.log2_min_pu_size; yy++)
            for (int xx = x >> p->sp.log2_min_pu_size; xx < (x + s) >> p->sp.log2_min_pu_size; xx++)
                f->is_pcm[yy * f->min_pu_width + xx] = 1;

    if (p->intra_pct && rnd_pct(&g->r, p->intra_pct)) {                  /* MODE_INTRA, hevcdec.c:2596-2640 */
        int nxn = log2 == 3 && rnd_pct(&g->r, 40), modes[4], mode_c = draw_mode(g);
        for (int i = 0; i < 4; i++) modes[i] = (nxn || !i) ? draw_mode(g) : modes[0];
        paint_bs(g, x, y, s, s);
        g->intra = 1;
        gen_tt_intra(g, x, y, log2, bypass, modes, mode_c, nxn, 0);
        g->intra = 0;
        return;
    }
    mark_done(g, x, y, s, s);
    mode = rnd_range(&g->r, 0, 99);
    if (log2 == 3) {
        if (mode < 60)      emit_pu(g, x, y, s, s);
        else if (mode < 80) { emit_pu(g, x, y, s, hlf); emit_pu(g, x, y + hlf, s, hlf); }
        else                { emit_pu(g, x, y, hlf, s); emit_pu(g, x + hlf, y, hlf, s); }
    } else {
        if (mode < 50)      emit_pu(g, x, y, s, s);                                               /* 2Nx2N */
        else if (mode < 62) { emit_pu(g, x, y, s, hlf); emit_pu(g, x, y + hlf, s, hlf); }         /* 2NxN  */
        else if (mode < 74) { emit_pu(g, x, y, hlf, s); emit_pu(g, x + hlf, y, hlf, s); }         /* Nx2N  */
        else if (mode < 81) { emit_pu(g, x, y, s, q); emit_pu(g, x, y + q, s, s - q); }           /* 2NxnU */
        else if (mode < 88) { emit_pu(g, x, y, s, s - q); emit_pu(g, x, y + s - q, s, q); }       /* 2NxnD */
        else if (mode < 94) { emit_pu(g, x, y, q, s); emit_pu(g, x + q, y, s - q, s); }           /* nLx2N */
        else                { emit_pu(g, x, y, s - q, s); emit_pu(g, x + s - q, y, q, s); }       /* nRx2N */
    }
    paint_bs(g, x, y, s, s);
    gen_tt(g, x, y, log2, bypass);
}

/* hls_coding_quadtree, hevcdec.c:2711 */
static void gen_cqt(gen *g, int x, int y, int log2)
{
    int s = 1 << log2;
    if (x >= g->W || y >= g->H) return;
    if (log2 > 3 && (x + s > g->W || y + s > g->H || rnd_pct(&g->r, g->p->split_pct[6 - log2]))) {
        int h = s >> 1;
        for (int i = 0; i < 4; i++)
            gen_cqt(g, x + (i & 1) * h, y + (i >> 1) * h, log2 - 1);
        return;
    }
    gen_cu(g, x, y, log2);
}

/* ------------------------------------------------------------------ frame */
void ohsynth_default_params(ohsynth_params *p, int width, int height, int bit_depth,
                            int chroma_format_idc)
{
    memset(p, 0, sizeof(*p));
    p->sp.width = width; p->sp.height = height; p->sp.bit_depth = bit_depth;
    p->sp.chroma_format_idc = chroma_format_idc;
    p->sp.log2_ctb_size = 6; p->sp.log2_min_cb_size = 3; p->sp.log2_min_pu_size = 2;
    p->sp.dpb_slots = 4; p->sp.jobs = 2; p->sp.sao_enabled = 1;
    p->sp.cb_qp_offset = 1; p->sp.cr_qp_offset = -2;
    p->seed = 1;
    p->n_slices = 1; p->tile_cols = 1; p->tile_rows = 1;
    p->far_mv_pct = 3; p->cbf_pct = 70; p->dc_pct = 15; p->dst_pct = 10; p->bi_pct = 60;
    p->split_pct[0] = 70; p->split_pct[1] = 50; p->split_pct[2] = 40;
    p->bs_pct[0] = 40; p->bs_pct[1] = 30;
    p->sao_pct[0] = 30; p->sao_pct[1] = 50;
    p->n_refs = 2;
}

static void gen_slice_tab(gen *g, ohcu_slice_tab *t)
{
    const ohsynth_params *p = g->p;
    memset(t, 0, sizeof(*t));
    for (int l = 0; l < 2; l++)
        for (int i = 0; i < OHCU_MAX_REFS; i++)
            t->ref_slot[l][i] = (uint8_t)((i + l) % p->n_refs);   /* L1 = L0 rotated */
    t->weight_flag = (uint8_t)!!p->weighted;
    t->luma_log2_weight_denom = (uint8_t)rnd_range(&g->r, 0, 7);
    t->chroma_log2_weight_denom = (uint8_t)rnd_range(&g->r, 0, 7);
    for (int l = 0; l < 2; l++)
        for (int i = 0; i < OHCU_MAX_REFS; i++) {
            /* pred_weight_table, hevcdec.c:202-306: w = (1<<denom) + delta, delta and
             * offsets in [-128,127] */
            int full = rnd_pct(&g->r, 15);      /* mostly mild weights so few samples saturate */
            int dl = full ? 128 : imin(128, ((1 << t->luma_log2_weight_denom) >> 1) + 1);
            int dc = full ? 128 : imin(128, ((1 << t->chroma_log2_weight_denom) >> 1) + 1);
            int ob = full ? 128 : 24;
            t->luma_weight[l][i] = (int16_t)((1 << t->luma_log2_weight_denom) + rnd_range(&g->r, -dl, dl - 1));
            t->luma_offset[l][i] = (int16_t)rnd_range(&g->r, -ob, ob - 1);
            for (int c = 0; c < 2; c++) {
                t->chroma_weight[l][i][c] = (int16_t)((1 << t->chroma_log2_weight_denom) + rnd_range(&g->r, -dc, dc - 1));
                t->chroma_offset[l][i][c] = (int16_t)rnd_range(&g->r, -ob, ob - 1);
            }
        }
}

static void gen_sao(gen *g, ohcu_sao *s)
{
    const ohsynth_params *p = g->p;
    int max_abs = (1 << (imin(g->bd, 10) - 5)) - 1;             /* hevc_cabac.c:712 */
    memset(s, 0, sizeof(*s));
    for (int c = 0; c < (g->cfi ? 3 : 1); c++) {
        if (c == 2) {                                           /* hevcdec.c:1261-1264 */
            s->type_idx[2] = s->type_idx[1];
            s->eo_class[2] = s->eo_class[1];
        } else {
            int v = rnd_range(&g->r, 0, 99);
            s->type_idx[c] = v < p->sao_pct[0] ? OHCU_SAO_BAND
                           : (v < p->sao_pct[0] + p->sao_pct[1] ? OHCU_SAO_EDGE : OHCU_SAO_NOT_APPLIED);
            s->eo_class[c] = (uint8_t)rnd_range(&g->r, 0, 3);
        }
        if (s->type_idx[c] == OHCU_SAO_NOT_APPLIED)
            continue;
        s->band_position[c] = (uint8_t)rnd_range(&g->r, 0, 31);
        s->offset_val[c][0] = 0;                                /* hevcdec.c:1289-1299 */
        for (int i = 0; i < 4; i++) {
            int v = rnd_range(&g->r, 0, max_abs);
            if (s->type_idx[c] == OHCU_SAO_EDGE) { if (i > 1) v = -v; }
            else if (rnd_pct(&g->r, 50)) v = -v;
            s->offset_val[c][i + 1] = (int16_t)v;
        }
    }
}

ohsynth_frame *ohsynth_generate(const ohsynth_params *p)
{
    gen *g = calloc(1, sizeof(*g));
    ohsynth_frame *f = calloc(1, sizeof(*f));
    const ohcu_seq_params *sp = &p->sp;
    int ctb = 1 << sp->log2_ctb_size, nctb;
    int *slice_start;

    g->p = p; g->f = f;
    g->W = sp->width; g->H = sp->height; g->bd = sp->bit_depth; g->cfi = sp->chroma_format_idc;
    g->hs = (g->cfi == 1 || g->cfi == 2); g->vs = (g->cfi == 1);
    g->qp = 30;
    rng_seed(&g->r, p->seed * 0x9E3779B97F4A7C15ull + 0x1234567);
    build_scan(g);

    f->bs_width = (sp->width >> 2) + 1;  f->bs_height = (sp->height >> 2) + 1;
    f->min_cb_width = sp->width >> sp->log2_min_cb_size; f->min_cb_height = sp->height >> sp->log2_min_cb_size;
    f->min_pu_width = sp->width >> sp->log2_min_pu_size; f->min_pu_height = sp->height >> sp->log2_min_pu_size;
    f->ctb_width = (sp->width + ctb - 1) >> sp->log2_ctb_size;
    f->ctb_height = (sp->height + ctb - 1) >> sp->log2_ctb_size;
    nctb = f->ctb_width * f->ctb_height;
    g->cell_w = (sp->width + 3) >> 2; g->cell_h = (sp->height + 3) >> 2;
    g->done = calloc((size_t)g->cell_w * g->cell_h, 1);
    f->vbs = calloc((size_t)f->bs_width * f->bs_height, 1);
    f->hbs = calloc((size_t)f->bs_width * f->bs_height, 1);
    f->qp_y_tab = calloc((size_t)f->min_cb_width * f->min_cb_height, 1);
    f->is_pcm = calloc((size_t)f->min_pu_width * f->min_pu_height, 1);
    f->deblock = calloc(nctb, sizeof(ohcu_dbparams));
    f->sao = calloc(nctb, sizeof(ohcu_sao));
    f->filter_slice_edges = calloc(nctb, 1);
    f->tab_slice_address = calloc(nctb, sizeof(int32_t));
    f->tile_id = calloc(nctb, sizeof(int32_t));
    /* at most one leaf per 4x4 luma block */
    f->leaves = calloc((size_t)((sp->width + 3) >> 2) * ((sp->height + 3) >> 2) + 1, sizeof(ohsynth_leaf));

    /* slices: n_slices-1 random cut points in CTB raster order */
    f->n_slices = imin(imin(p->n_slices, OHCU_MAX_SLICES), nctb);
    f->slices = calloc(f->n_slices, sizeof(ohcu_slice_tab));
    slice_start = calloc(f->n_slices + 1, sizeof(int));
    for (int i = 1; i < f->n_slices; i++)
        slice_start[i] = iclip((int)((int64_t)nctb * i / f->n_slices) + rnd_range(&g->r, -2, 2), slice_start[i - 1] + 1, nctb - (f->n_slices - i));
    slice_start[f->n_slices] = nctb;
    for (int i = 0; i < f->n_slices; i++)
        gen_slice_tab(g, &f->slices[i]);

    for (int rs = 0, si = 0; rs < nctb; rs++) {
        int cx = rs % f->ctb_width, cy = rs / f->ctb_width;
        while (rs >= slice_start[si + 1]) si++;
        g->slice = si;
        f->tab_slice_address[rs] = slice_start[si];
        f->filter_slice_edges[rs] = (uint8_t)!(p->slices_no_lf && (si & 1));
        f->tile_id[rs] = (cy * p->tile_rows / f->ctb_height) * p->tile_cols + cx * p->tile_cols / f->ctb_width;
        /* slice_beta_offset / slice_tc_offset are *_div2 * 2 in [-12,12]; drawn per CTB so
         * that the left-CTB / current-CTB offset rules of hevc_filter.c:666-721 are hit */
        f->deblock[rs].beta_offset = (int8_t)(2 * rnd_range(&g->r, -6, 6));
        f->deblock[rs].tc_offset = (int8_t)(2 * rnd_range(&g->r, -6, 6));
        gen_sao(g, &f->sao[rs]);
        gen_cqt(g, cx << sp->log2_ctb_size, cy << sp->log2_ctb_size, sp->log2_ctb_size);
    }
    free(slice_start);
    free(g->done);
    free(g);
    return f;
}

void ohsynth_free(ohsynth_frame *f)
{
    if (!f) return;
    free(f->slices); free(f->pus); free(f->tus); free(f->coeffs); free(f->vbs); free(f->hbs);
    free(f->qp_y_tab); free(f->deblock); free(f->is_pcm); free(f->sao);
    free(f->filter_slice_edges); free(f->tab_slice_address); free(f->tile_id); free(f->leaves); free(f->ius);
    free(f);
}

void ohsynth_fill_plane(uint8_t *plane, ptrdiff_t stride, int w, int h, int bit_depth, uint64_t seed)
{
    rng r;
    int maxv = (1 << bit_depth) - 1;
    rng_seed(&r, seed ^ 0xA5A5A5A5DEADBEEFull);
    int gx = rnd_range(&r, 1, 5), gy = rnd_range(&r, 1, 5), g0 = rnd_range(&r, 0, maxv);
    for (int y = 0; y < h; y++) {
        uint64_t bits = 0; int nb = 0;
        for (int x = 0; x < w; x++) {
            int noise, grad, v;
            if (nb < 16) { bits = rnd(&r); nb = 64; }
            noise = (int)(bits & 0xFFFF) & maxv; bits >>= 16; nb -= 16;
            grad = (g0 + ((x * gx + y * gy) << (bit_depth - 8)) / 4) & (2 * maxv + 1);
            if (grad > maxv) grad = 2 * maxv + 1 - grad;        /* triangle wave */
            /* 3/4 of the samples are a smooth ramp + small noise (so that deblocking
             * decisions take every branch), 1/4 are full-range noise */
            if ((x >> 4 ^ y >> 4) & 3)
                v = iclip(grad + (noise >> (bit_depth - 3)) - 4, 0, maxv);
            else
                v = noise;
            if (bit_depth == 8) plane[y * stride + x] = (uint8_t)v;
            else ((uint16_t *)(plane + y * stride))[x] = (uint16_t)v;
        }
    }
}
