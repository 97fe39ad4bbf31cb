/*
 * e2e_driver.c — a plain-C, multi-threaded caller of include/ohcu.h: the host side of a decoder reduced
 * to what this path sees.  Frame threads (the reference runs up to 16, hevcdec.h:53, each with its own
 * HEVCLocalContext, hevcdec.h:508-510) take frames from a counter; for each one a thread begins a job, writes
 * the frame's work lists into the job's pinned arenas the way the patched call sites of hevcdec.c /
 * hevc_cabac.c / hevc_filter.c would, commits, flushes and asks for the finished picture.
 *
 * It is the measurement harness of bench.py's e2e leg (no Python, no GIL, no ctypes marshalling inside
 * the timed region) and the first C program that binds the ABI (tests/test_gpu_driver.py runs it and
 * compares every picture with the oracle).  Built as hevc_b200/libohcu_drv.so next to libohcu.so.
 *
 * A "frame" here is a set of lists prepared ahead (synthetic, hevc_b200/synth): standing in for the parser,
 * the thread memcpy's them into the arenas inside the timed region — a lower bound of the stores a real
 * parser makes while it decodes (OHDRV_FILL_COPY) — or finds the big lists already there (OHDRV_FILL_NONE,
 * the library alone).  OHDRV_APPEND drives the copying ohcu_append_* calls instead.
 */
#define _GNU_SOURCE
#include <pthread.h>
#include <sched.h>
#include <stdatomic.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <time.h>

#include <stdlib.h>

#include "e2e_driver.h"

/* OHDRV_TRACE=1: one line per step on stderr (debugging a stuck run) */
static int g_trace = -1;
static double now_ms(clockid_t id)
{
    struct timespec ts;
    clock_gettime(id, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}
#define TRACE(...) do { if (g_trace < 0) g_trace = getenv("OHDRV_TRACE") != NULL; if (g_trace) { \
    fprintf(stderr, "[%.1f] ", now_ms(CLOCK_MONOTONIC)); fprintf(stderr, __VA_ARGS__); fflush(stderr); } } while (0)

typedef struct run_state {
    ohcu_ctx *ctx;
    const ohdrv_frame *frames;
    int n_frames;
    const ohdrv_config *cfg;
    atomic_int next;
    atomic_int failed;
    atomic_int turn[64];                 /* per ring entry (or per stream): which of its frames may be submitted next */
    ohcu_seq_params sp;
    int nctb;
    size_t bs_bytes, qp_bytes, pcm_bytes;
    char err[512];
    pthread_mutex_t err_mu;
} run_state;

static void set_err(run_state *st, const char *what, int rc)
{
    pthread_mutex_lock(&st->err_mu);
    if (!atomic_load(&st->failed)) {
        snprintf(st->err, sizeof(st->err), "%s -> %d: %s", what, rc, ohcu_last_error());
        atomic_store(&st->failed, 1);
    }
    pthread_mutex_unlock(&st->err_mu);
}

/* the frame's lists into the job's arenas, then the counts */
static int fill_inplace(run_state *st, int job, const ohdrv_frame *f, int copy, int slot_a, int slot_b)
{
    ohcu_frame_arenas a;
    ohcu_frame_counts n;
    int rc = ohcu_frame_arenas_get(st->ctx, job, &a);
    if (rc) return rc;
    memset(&n, 0, sizeof(n));
    if (copy) {
        memcpy(a.pus, f->pus, (size_t)f->n_pu * sizeof(ohcu_pu));
        memcpy(a.tus, f->tus, (size_t)f->n_tu * sizeof(ohcu_tu));
        if (f->coef_form == OHCU_COEF_LEVELS) memcpy(a.quant, f->quant, (size_t)f->n_tu * sizeof(ohcu_tu_quant));
        memcpy(a.coefs, f->coefs, (size_t)f->n_coef * (f->coef_form >= OHCU_COEF_SPARSE ? sizeof(ohcu_coef) : sizeof(int16_t)));
        if (f->n_iu) memcpy(a.ius, f->ius, (size_t)f->n_iu * sizeof(ohcu_iu));
    }
    /* the small tables are written every frame in both modes */
    memcpy(a.ctbs, f->ctbs, (size_t)st->nctb * sizeof(ohcu_ctb_index));
    if (f->n_iu) {
        if (!a.ius) return OHCU_ESTATE;
        memcpy(a.iu_ctbs, f->iu_ctbs, (size_t)st->nctb * sizeof(ohcu_iu_index));
    }
    memcpy(a.slices, f->slices, (size_t)f->n_slices * sizeof(ohcu_slice_tab));
    if (slot_a >= 0)                                   /* chain mode: this picture predicts from the two previous ones */
        for (int s = 0; s < f->n_slices; s++)
            for (int l = 0; l < 2; l++)
                for (int i = 0; i < OHCU_MAX_REFS; i++) a.slices[s].ref_slot[l][i] = (uint8_t)(((i + l) & 1) ? slot_b : slot_a);
    if (f->scaling_list) { memcpy(a.scaling_list, f->scaling_list, 4 * 6 * 64 + 2 * 6); n.scaling_list = 1; }
    if (f->deblock) {
        memcpy(a.qp_y_tab, f->qp, st->qp_bytes);
        memcpy(a.deblock, f->db, (size_t)st->nctb * sizeof(ohcu_dbparams));
        if (f->deblock == OHCU_DEBLOCK_PLANES) {
            memcpy(a.vertical_bs, f->vbs, st->bs_bytes);
            memcpy(a.horizontal_bs, f->hbs, st->bs_bytes);
        } else {
            memset(a.deblock_disabled, 0, (size_t)st->nctb);
        }
        if (f->is_pcm) { memcpy(a.is_pcm, f->is_pcm, st->pcm_bytes); n.pcm = 1; }
    }
    if (f->has_sao || f->deblock == OHCU_DEBLOCK_DERIVED) {
        memcpy(a.filter_slice_edges, f->fse, (size_t)st->nctb);
        memcpy(a.tab_slice_address, f->tsa, (size_t)st->nctb * sizeof(int32_t));
        memcpy(a.tile_id, f->tile, (size_t)st->nctb * sizeof(int32_t));
    }
    if (f->has_sao) memcpy(a.sao, f->sao, (size_t)st->nctb * sizeof(ohcu_sao));
    n.n_pu = f->n_pu; n.n_tu = f->n_tu; n.n_coef = f->n_coef; n.n_slices = f->n_slices; n.coef_form = f->n_tu ? f->coef_form : 0;
    n.deblock = f->deblock; n.sao = f->has_sao; n.all_inter = f->all_inter; n.n_iu = f->n_iu;
    return ohcu_frame_commit_lists(st->ctx, job, &n);
}

/* the same frame through the copying entry points */
static int fill_append(run_state *st, int job, const ohdrv_frame *f)
{
    int rc;
    for (int s = 0; s < f->n_slices; s++)
        if ((rc = ohcu_set_slice(st->ctx, job, s, &f->slices[s]))) return rc;
    if ((rc = ohcu_append_pu(st->ctx, job, f->pus, (int)f->n_pu))) return rc;
    if (f->n_iu && (rc = ohcu_append_iu(st->ctx, job, f->ius, (int)f->n_iu))) return rc;
    if (f->n_tu) {
        if (f->coef_form == OHCU_COEF_LEVELS) {
            if (f->scaling_list &&
                (rc = ohcu_set_scaling_list(st->ctx, job, (const uint8_t(*)[6][64])f->scaling_list,
                                            (const uint8_t(*)[6])(f->scaling_list + 4 * 6 * 64)))) return rc;
            rc = ohcu_append_tu_levels(st->ctx, job, f->tus, f->quant, (int)f->n_tu, (const ohcu_coef *)f->coefs, (size_t)f->n_coef);
        } else if (f->coef_form == OHCU_COEF_SPARSE) {
            rc = ohcu_append_tu_sparse(st->ctx, job, f->tus, (int)f->n_tu, (const ohcu_coef *)f->coefs, (size_t)f->n_coef);
        } else {
            rc = ohcu_append_tu(st->ctx, job, f->tus, (int)f->n_tu, (const int16_t *)f->coefs, (size_t)f->n_coef);
        }
        if (rc) return rc;
    }
    if (f->deblock == OHCU_DEBLOCK_PLANES) {
        ohcu_deblock_in d = { f->vbs, f->hbs, f->qp, f->db, f->is_pcm };
        if ((rc = ohcu_set_deblock(st->ctx, job, &d))) return rc;
    } else if (f->deblock == OHCU_DEBLOCK_DERIVED) {
        ohcu_deblock_derive_in d = { f->qp, f->db, f->is_pcm, f->fse, f->tsa, f->tile, NULL };
        if ((rc = ohcu_set_deblock_derived(st->ctx, job, &d))) return rc;
    }
    if (f->has_sao) {
        ohcu_sao_in s = { f->sao, f->fse, f->tsa, f->tile };
        if ((rc = ohcu_set_sao(st->ctx, job, &s))) return rc;
    }
    return 0;
}

typedef struct thread_arg { run_state *st; double cpu_ms, work_ms; int frames; } thread_arg;

/* one frame: wait for its turn on its ring entry / stream, begin, fill, flush, ask for the picture */
static int submit_frame(run_state *st, int i, double *work_ms)
{
    const ohdrv_config *cfg = st->cfg;
    int k, key, ticket, slot_a = -1, slot_b = -1, fidx, rc = 0;
    if (cfg->chain_streams > 0) {
        /* S independent streams, each a chain: picture t of stream s goes to entry t mod R of the stream's R ring
         * entries and predicts from the stream's two previous pictures (what a decoder does) */
        const int S = cfg->chain_streams, R = cfg->ring / S, s = i % S, t = i / S;
        k = s * R + t % R;
        slot_a = t >= 1 ? cfg->first_slot + s * R + (t - 1) % R : cfg->ref_slot[0];
        slot_b = t >= 2 ? cfg->first_slot + s * R + (t - 2) % R : cfg->ref_slot[1];
        key = s; ticket = t; fidx = t;
    } else {
        k = i % cfg->ring;
        key = k; ticket = i / cfg->ring; fidx = i;
    }
    const int out_slot = cfg->first_slot + k;
    /* frames that share a ring entry (their output slot and pinned picture), or a stream, are submitted in order */
    while (atomic_load(&st->turn[key]) != ticket) {
        if (atomic_load(&st->failed)) return -1;
        sched_yield();
    }
    TRACE("frame %d: turn ok (key %d ticket %d), begin slot %d\n", i, key, ticket, out_slot);
    int job;
    for (;;) {                               /* OHCU_ESTATE: the ring's next job is still being filled by a slower thread */
        job = ohcu_frame_begin(st->ctx, out_slot, i);
        if (job != OHCU_ESTATE || atomic_load(&st->failed)) break;
        sched_yield();
    }
    if (job < 0) { set_err(st, "ohcu_frame_begin", job); rc = job; goto out; }
    TRACE("frame %d: job %d\n", i, job);
    const double w0 = now_ms(CLOCK_THREAD_CPUTIME_ID);           /* the library's and the fill's own CPU time: no waiting in it */
    if (cfg->frame_by_job) fidx = job;
    const ohdrv_frame *f = &st->frames[fidx % st->n_frames];
    if (cfg->fill == OHDRV_APPEND) rc = fill_append(st, job, f);
    else rc = fill_inplace(st, job, f, cfg->fill == OHDRV_FILL_COPY, slot_a, slot_b);
    if (rc < 0) { set_err(st, "filling the job's lists", rc); ohcu_frame_abort(st->ctx, job); goto out; }
    TRACE("frame %d: filled\n", i);
    if ((rc = ohcu_frame_flush(st->ctx, job)) < 0) { set_err(st, "ohcu_frame_flush", rc); goto out; }
    TRACE("frame %d: flushed\n", i);
    if (cfg->download == OHDRV_DOWNLOAD_PLANES) {
        if ((rc = ohcu_slot_download_async(st->ctx, out_slot, cfg->out_planes + 3 * k, cfg->out_strides)) < 0)
            set_err(st, "ohcu_slot_download_async", rc);
    } else if (cfg->download == OHDRV_DOWNLOAD_CROPPED) {
        ohcu_crop crop = { 0, 0, st->sp.width, st->sp.height };
        if ((rc = ohcu_slot_output_cropped_async(st->ctx, out_slot, &crop, cfg->out_planes + 3 * k)) < 0)
            set_err(st, "ohcu_slot_output_cropped_async", rc);
    }
    if (cfg->last_frame) cfg->last_frame[k] = fidx % st->n_frames;
    *work_ms += now_ms(CLOCK_THREAD_CPUTIME_ID) - w0;
    TRACE("frame %d: submitted\n", i);
out:
    atomic_store(&st->turn[key], ticket + 1);
    return rc;
}

static void *frame_thread(void *p)
{
    thread_arg *ta = (thread_arg *)p;
    run_state *st = ta->st;
    const double t0 = now_ms(CLOCK_THREAD_CPUTIME_ID);
    for (;;) {
        const int i = atomic_fetch_add(&st->next, 1);
        if (i >= st->cfg->total_frames) break;
        if (atomic_load(&st->failed)) {                    /* keep the tickets moving so that nobody waits for ever */
            const ohdrv_config *cfg = st->cfg;
            const int key = cfg->chain_streams > 0 ? i % cfg->chain_streams : i % cfg->ring;
            const int ticket = cfg->chain_streams > 0 ? i / cfg->chain_streams : i / cfg->ring;
            while (atomic_load(&st->turn[key]) != ticket) sched_yield();
            atomic_store(&st->turn[key], ticket + 1);
            continue;
        }
        if (submit_frame(st, i, &ta->work_ms) >= 0) ta->frames++;
    }
    ta->cpu_ms = now_ms(CLOCK_THREAD_CPUTIME_ID) - t0;
    return NULL;
}

int ohdrv_run(ohcu_ctx *ctx, const ohdrv_frame *frames, int n_frames, const ohdrv_config *cfg, ohdrv_result *res)
{
    static run_state st;                 /* one run at a time per process (the harness is single-purpose) */
    static pthread_mutex_t run_mu = PTHREAD_MUTEX_INITIALIZER;
    pthread_t th[OHDRV_MAX_THREADS];
    thread_arg ta[OHDRV_MAX_THREADS];
    if (!ctx || !frames || !cfg || !res || n_frames < 1 || cfg->threads < 1 || cfg->threads > OHDRV_MAX_THREADS ||
        cfg->ring < 1 || cfg->ring > 64 || cfg->total_frames < 0 ||
        (cfg->chain_streams > 0 && (cfg->ring % cfg->chain_streams || cfg->ring / cfg->chain_streams < 3)) ||
        (cfg->download != OHDRV_DOWNLOAD_NONE && !cfg->out_planes) || (cfg->download == OHDRV_DOWNLOAD_PLANES && !cfg->out_strides))
        return OHCU_EINVAL;
    pthread_mutex_lock(&run_mu);
    memset(&st, 0, sizeof(st));
    memset(res, 0, sizeof(*res));
    st.ctx = ctx; st.frames = frames; st.n_frames = n_frames; st.cfg = cfg;
    int rc = ohcu_get_seq_params(ctx, &st.sp);
    if (rc) { pthread_mutex_unlock(&run_mu); return rc; }
    {
        const int cs = 1 << st.sp.log2_ctb_size;
        st.nctb = ((st.sp.width + cs - 1) / cs) * ((st.sp.height + cs - 1) / cs);
        st.bs_bytes = (size_t)((st.sp.width >> 2) + 1) * ((st.sp.height >> 2) + 1);
        st.qp_bytes = (size_t)(st.sp.width >> st.sp.log2_min_cb_size) * (st.sp.height >> st.sp.log2_min_cb_size);
        st.pcm_bytes = (size_t)(st.sp.width >> st.sp.log2_min_pu_size) * (st.sp.height >> st.sp.log2_min_pu_size);
    }
    atomic_init(&st.next, 0);
    atomic_init(&st.failed, 0);
    for (int i = 0; i < 64; i++) atomic_init(&st.turn[i], 0);
    pthread_mutex_init(&st.err_mu, NULL);
    if (cfg->timed && (rc = ohcu_timer_start(ctx))) { pthread_mutex_unlock(&run_mu); return rc; }
    const double w0 = now_ms(CLOCK_MONOTONIC);
    for (int t = 0; t < cfg->threads; t++) {
        ta[t].st = &st; ta[t].cpu_ms = 0; ta[t].work_ms = 0; ta[t].frames = 0;
        pthread_create(&th[t], NULL, frame_thread, &ta[t]);
    }
    for (int t = 0; t < cfg->threads; t++) pthread_join(th[t], NULL);
    TRACE("threads joined\n");
    res->host_ms = now_ms(CLOCK_MONOTONIC) - w0;
    if (cfg->timed) rc = ohcu_timer_stop(ctx, &res->device_ms);
    else rc = ohcu_sync_all(ctx);
    res->wall_ms = now_ms(CLOCK_MONOTONIC) - w0;
    for (int t = 0; t < cfg->threads; t++) { res->thread_cpu_ms += ta[t].cpu_ms; res->work_cpu_ms += ta[t].work_ms; res->frames += ta[t].frames; }
    pthread_mutex_destroy(&st.err_mu);
    if (atomic_load(&st.failed)) {
        snprintf(res->error, sizeof(res->error), "%s", st.err);
        rc = OHCU_EINVAL;
    }
    /* list errors found on the device surface at the slots' sync */
    for (int k = 0; k < cfg->ring && k < cfg->total_frames && !rc; k++) {
        rc = ohcu_slot_sync(ctx, cfg->first_slot + k);
        if (rc) snprintf(res->error, sizeof(res->error), "ohcu_slot_sync(%d) -> %d: %s", cfg->first_slot + k, rc, ohcu_last_error());
    }
    pthread_mutex_unlock(&run_mu);
    return rc;
}
