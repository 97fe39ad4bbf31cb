// api.cu — the C ABI of include/ohcu.h: context, device DPB, per-frame job arenas,
// list staging and the launch sequence.  Host-side counterpart of what set_sps /
// pic_arrays_init / hevc_frame_start / the hevc_refs.c frame pool do for this path in
// the reference (hevcdec.c:114-200, 428-476, 3575-3600; hevc_refs.c:79-129).
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <new>
#include <utility>
#include <vector>

#include "ohcu_internal.h"

using namespace ohcu;

// ------------------------------------------------------------------ errors
static thread_local char g_err[512] = "";
static int fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
// the same for entry points that live beside their kernels (recon.cu: ohcu_dsp_idct_noalloc)
namespace ohcu {
int set_error(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
}  // namespace ohcu
#define CU(call)                                                                          \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess)                                                            \
            return fail(OHCU_ECUDA, "%s -> %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// ------------------------------------------------------------------ objects
struct Picture {
    uint8_t *base = nullptr;
    uint8_t *padded_base[3] = {nullptr, nullptr, nullptr};   // first byte of each (padded) plane
    PlaneSet ps{};
    size_t bytes = 0;
};

struct Slot {
    Picture pic;
    cudaEvent_t ready = nullptr;      // producer finished
    cudaEvent_t read_done = nullptr;  // last async download finished
    bool has_ready = false, has_read = false;
    uint64_t readers = 0;             // jobs that read this slot since its last write
    uint8_t *out_d = nullptr;         // tight staging of ohcu_slot_output_cropped (allocated on first use)
    int producer = -1;                // job that wrote the slot last (its error word is the slot's)
    bool released = false;            // ohcu_slot_release: no picture to read until the slot is written again
    int copy_lane = 0;                // which copy stream the slot's downloads use
};

template <typename T> struct Arena {      // pinned host staging + device copy
    T *h = nullptr;
    T *d = nullptr;
    size_t cap = 0;
    bool carved = false;                  // lives inside the job's table block (Job::tab_h / tab_d)
};

constexpr size_t SL_BYTES = 4 * 6 * 64 + 2 * 6;      // ScalingList, hevc_ps.h:605-610
constexpr int COPY_LANES = 4;

struct Job {
    cudaStream_t stream = nullptr;
    cudaStream_t side = nullptr;      // the bs derivation runs beside mc_kernel and residual_kernel (it needs only the lists)
    cudaEvent_t fork = nullptr, join = nullptr;
    cudaEvent_t done = nullptr;
    cudaEvent_t ev[OHCU_NUM_STAGES + 1] = {};
    bool has_done = false, timed = false;
    bool open = false;                // between frame_begin and stage
    bool staged = false;
    int out_slot = -1, poc = 0;
    Picture recon;
    LfMaps lf_maps{};                 // recon planes as the loop filter kernel fetches them
    // lists
    Arena<ohcu_pu> pus;      int n_pu = 0;
    Arena<int2> ctb_range;   // per CTB: first PU, count
    bool pu_sorted = true;   int last_ctb = -1;
    Arena<ohcu_slice_tab> slices; int n_slices = 0;
    std::vector<int> ctb_area;   // luma samples covered by PUs, per CTB
    bool all_covered = false;
    Arena<ohcu_tu> tus;      int n_tu = 0;
    Arena<int2> ctb_tu_range;    // per CTB: first TU, count
    bool tu_sorted = true;   int last_tu_ctb = -1;
    bool has_cross = false;  int max_slice = -1;    // what frame_stage has to validate
    Arena<int16_t> coeffs;   size_t ncoeff = 0;
    uint8_t *tab_h = nullptr, *tab_d = nullptr;      // block the table arenas are carved from
    size_t tab_end[5] = {0, 0, 0, 0, 0};             // bytes up to and including: CTB ranges + slices, SAO, qp + DBParams + bs flags, bs planes, is_pcm
    Arena<uint8_t> bs_flags;     // per CTB BSF_*, bs derived on the device
    bool bs_derived = false;
    // sparse form (ohcu_append_tu_sparse): the arena holds ohcu_coef pairs, ncoeff counts int16
    Arena<int2> ctb_coef_range;  // per CTB: first pair, number of pairs
    int coef_form = 0;           // 0 none yet, 1 dense, 2 sparse, 3 sparse un-scaled levels + per-TU quantiser
    Arena<ohcu_tu_quant> tuq;    // levels form: parallel to tus
    uint8_t *sl_h = nullptr, *sl_d = nullptr;   // ScalingList (sl[4][6][64], sl_dc[2][6])
    bool has_sl = false;
    std::vector<uint32_t> ctb_seen; uint32_t seen_epoch = 0;   // CTBs met by the current sparse append
    Arena<uint8_t> vbs, hbs, is_pcm; Arena<int8_t> qp; Arena<ohcu_dbparams> db;
    Arena<ohcu_sao> sao;     Arena<uint16_t> sao_flags;
    bool has_deblock = false, has_sao = false, has_pcm = false;
    ohcu_pic_params pp{};             // picture-level PPS fields of this frame (default: the context's)
    // in-place lists (ohcu_frame_arenas_get / ohcu_frame_commit_lists): the parser wrote into the pinned arenas,
    // prepare_kernel validates and indexes them on the device
    Arena<ohcu_ctb_index> ctbs;
    Arena<uint8_t> fse, dbk_disabled; Arena<int32_t> tsa, tile;
    bool inplace = false; int all_inter = 0;
    uint64_t n_coef_units = 0;        // pairs (sparse / levels) or int16 (dense) committed
    int *err_d = nullptr, *err_h = nullptr;      // the job's error word (device) and its copy (pinned)
    // PCM blocks (ohcu_append_pcm): allocated on first use
    PcmRec *pcm_h = nullptr, *pcm_d = nullptr; int n_pcm = 0, pcm_cap = 0;
    uint8_t *pcms_h = nullptr, *pcms_d = nullptr; size_t pcm_bytes = 0, pcms_cap = 0;
    // intra blocks (ohcu_seq_params.intra): records, where each CTB's sit, the residual planes of their coded TBs,
    // the ticket and per-CTB done flags of intra_kernel
    Arena<ohcu_iu> ius;      int n_iu = 0;
    Arena<ohcu_iu_index> iu_ctbs; int last_iu_ctb = -1;
    uint8_t *resid_base = nullptr; PlaneSet resid{};
    int *progress_d = nullptr;
    bool inplace_iu = false;          // iu_ctbs was handed to the parser (zeroed again by the next ohcu_frame_begin)
};

struct ohcu_ctx {
    int device = 0;
    ohcu_seq_params sp{};
    int hs = 0, vs = 0, nplanes = 1, P = 1;
    int pw[3] = {0, 0, 0}, ph[3] = {0, 0, 0};
    int ctb_w = 0, ctb_h = 0, nctb = 0;
    int min_cb_w = 0, min_cb_h = 0, min_pu_w = 0, min_pu_h = 0, bs_w = 0, bs_h = 0;
    std::vector<Slot> slots;
    std::vector<Job> jobs;
    int next_job = 0;
    cudaStream_t copy_stream = nullptr;
    cudaStream_t copy_lanes[COPY_LANES] = {};    // downloads of different slots overlap (lane 0 = copy_stream)
    bool profiling = false;
    int64_t launches = 0;
    void *l2_buf = nullptr;
    size_t l2_bytes = 0;
    uint8_t *md5_d = nullptr;         // 48 bytes per slot (ohcu_slot_md5)
    int *intra_order_d = nullptr;     // CTB addresses in wavefront order (intra_kernel)
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    // Frame threads each drive their own job; the lists of a job are touched by one thread
    // at a time (caller's contract).  What jobs share — the job ring, the slots' ordering
    // events and reader masks, the launch counter — is guarded here.
    std::mutex mu;
};

// ------------------------------------------------------------------ helpers
// padded: DPB pictures carry the replicated halo the MC windows are fetched from
// (PAD_X / PAD_Y, ohcu_internal.h); work pictures do not.
static int alloc_picture(const ohcu_ctx *c, Picture *p, bool padded)
{
    const int px = padded ? PAD_X : 0, py = padded ? PAD_Y : 0;
    size_t off[3], total = 0;
    for (int i = 0; i < c->nplanes; i++) {
        p->ps.pitch[i] = ((c->pw[i] + 2 * px) * c->P + 255) & ~255;
        off[i] = total;
        total += (size_t)p->ps.pitch[i] * (c->ph[i] + 2 * py) + 256;
    }
    CU(cudaMalloc(&p->base, total));
    CU(cudaMemset(p->base, 0, total));
    for (int i = 0; i < 3; i++) {
        p->padded_base[i] = i < c->nplanes ? p->base + off[i] : nullptr;
        p->ps.p[i] = i < c->nplanes ? p->base + off[i] + (size_t)py * p->ps.pitch[i] + (size_t)px * c->P : nullptr;
    }
    p->bytes = total;
    return 0;
}

// The loop filter fetches its tiles with cp.async.bulk.tensor.2d through one tensor map per
// (job, plane), passed as kernel parameters.  The driver entry point is looked up at run time
// so that the library has no link-time dependency on libcuda.
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static void fill_seq_dev(const ohcu_ctx *c, FrameDev *f);
static int build_tensor_maps(ohcu_ctx *c)
{
    EncodeTiledFn enc = nullptr;
    cudaDriverEntryPointQueryResult qr;
    CU(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&enc, cudaEnableDefault, &qr));
    if (!enc || qr != cudaDriverEntryPointSuccess) return fail(OHCU_ECUDA, "cuTensorMapEncodeTiled is not available");
    const bool high = c->P == 2;
    // the loop filter reads each job's reconstructed picture through kernel-parameter maps
    FrameDev fbox{};
    fill_seq_dev(c, &fbox);
    for (auto &j : c->jobs)
        for (int pl = 0; pl < c->nplanes; pl++) {
            int bw, bh;
            loopfilter_box(fbox, pl, &bw, &bh);
            cuuint64_t gdim[2] = {(cuuint64_t)c->pw[pl], (cuuint64_t)c->ph[pl]};
            cuuint64_t gstr[1] = {(cuuint64_t)j.recon.ps.pitch[pl]};
            cuuint32_t box[2] = {(cuuint32_t)bw, (cuuint32_t)bh};
            cuuint32_t estr[2] = {1, 1};
            CUresult r = enc(&j.lf_maps.m[pl], high ? CU_TENSOR_MAP_DATA_TYPE_UINT16 : CU_TENSOR_MAP_DATA_TYPE_UINT8, 2,
                             j.recon.ps.p[pl], gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) return fail(OHCU_ECUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
        }
    return 0;
}

template <typename T> static int alloc_arena(Arena<T> *a, size_t n)
{
    a->cap = n ? n : 1;
    CU(cudaHostAlloc((void **)&a->h, a->cap * sizeof(T), cudaHostAllocDefault));
    CU(cudaMalloc((void **)&a->d, a->cap * sizeof(T)));
    return 0;
}
// The per-CTB / per-edge tables of a frame are carved out of one pinned block and one device
// block in upload order, so that a frame's tables go up in ONE copy: 13 separate copies per
// frame cost 11% of the link when uploads and downloads overlap (tools/pcie_probe3.py).
template <typename T> static void carve_arena(Arena<T> *a, size_t n, uint8_t *hbase, uint8_t *dbase, size_t *off)
{
    a->cap = n ? n : 1;
    a->carved = true;
    if (hbase) { a->h = (T *)(hbase + *off); a->d = (T *)(dbase + *off); }
    *off += (a->cap * sizeof(T) + 255) & ~(size_t)255;
}
template <typename T> static void free_arena(Arena<T> *a)
{
    if (a->h && !a->carved) cudaFreeHost(a->h);
    if (a->d && !a->carved) cudaFree(a->d);
    a->h = nullptr; a->d = nullptr; a->cap = 0;
}
template <typename T> static cudaError_t push_arena(const Arena<T> &a, size_t n, cudaStream_t st)
{
    if (!n) return cudaSuccess;
    return cudaMemcpyAsync(a.d, a.h, n * sizeof(T), cudaMemcpyHostToDevice, st);
}

static int check_job(ohcu_ctx *c, int job)
{
    if (!c) return fail(OHCU_EINVAL, "null context");
    if (job < 0 || job >= (int)c->jobs.size()) return fail(OHCU_EINVAL, "bad job id %d", job);
    return 0;
}
#define ENTER(c)                                              \
    do {                                                      \
        if (!(c)) return fail(OHCU_EINVAL, "null context");   \
        CU(cudaSetDevice((c)->device));                       \
    } while (0)

// ------------------------------------------------------------------ library
extern "C" int ohcu_abi_version(void) { return OHCU_ABI_VERSION; }
extern "C" const char *ohcu_last_error(void) { return g_err; }
extern "C" int ohcu_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

extern "C" int ohcu_destroy(ohcu_ctx *c)
{
    if (!c) return 0;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    for (auto &s : c->slots) {
        if (s.pic.base) cudaFree(s.pic.base);
        if (s.out_d) cudaFree(s.out_d);
        if (s.ready) cudaEventDestroy(s.ready);
        if (s.read_done) cudaEventDestroy(s.read_done);
    }
    for (auto &j : c->jobs) {
        if (j.recon.base) cudaFree(j.recon.base);
        free_arena(&j.pus); free_arena(&j.ctb_range); free_arena(&j.slices); free_arena(&j.tus);
        free_arena(&j.coeffs); free_arena(&j.ctb_tu_range); free_arena(&j.ctb_coef_range);
        free_arena(&j.vbs); free_arena(&j.hbs); free_arena(&j.is_pcm); free_arena(&j.qp); free_arena(&j.db);
        free_arena(&j.sao); free_arena(&j.sao_flags); free_arena(&j.tuq);
        if (j.sl_h) cudaFreeHost(j.sl_h);
        if (j.sl_d) cudaFree(j.sl_d);
        if (j.err_h) cudaFreeHost(j.err_h);
        if (j.err_d) cudaFree(j.err_d);
        if (j.pcm_h) cudaFreeHost(j.pcm_h);
        if (j.pcm_d) cudaFree(j.pcm_d);
        if (j.pcms_h) cudaFreeHost(j.pcms_h);
        if (j.pcms_d) cudaFree(j.pcms_d);
        free_arena(&j.ius); free_arena(&j.iu_ctbs);
        if (j.resid_base) cudaFree(j.resid_base);
        if (j.progress_d) cudaFree(j.progress_d);
        if (j.tab_h) cudaFreeHost(j.tab_h);
        if (j.tab_d) cudaFree(j.tab_d);
        if (j.done) cudaEventDestroy(j.done);
        for (auto &e : j.ev) if (e) cudaEventDestroy(e);
        if (j.fork) cudaEventDestroy(j.fork);
        if (j.join) cudaEventDestroy(j.join);
        if (j.side) cudaStreamDestroy(j.side);
        if (j.stream) cudaStreamDestroy(j.stream);
    }
    for (int i = 1; i < COPY_LANES; i++) if (c->copy_lanes[i]) cudaStreamDestroy(c->copy_lanes[i]);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->l2_buf) cudaFree(c->l2_buf);
    if (c->md5_d) cudaFree(c->md5_d);
    if (c->intra_order_d) cudaFree(c->intra_order_d);
    if (c->t0) cudaEventDestroy(c->t0);
    if (c->t1) cudaEventDestroy(c->t1);
    delete c;
    return 0;
}

extern "C" int ohcu_create(ohcu_ctx **out, int device, const ohcu_seq_params *sp)
{
    if (!out || !sp) return fail(OHCU_EINVAL, "null argument");
    *out = nullptr;
    if (sp->width <= 0 || sp->height <= 0 || sp->width > 16384 || sp->height > 16384)
        return fail(OHCU_EINVAL, "bad picture size %dx%d", sp->width, sp->height);
    // the depths ff_hevc_dsp_init instantiates (hevcdsp.c:514-545)
    if (sp->bit_depth != 8 && sp->bit_depth != 9 && sp->bit_depth != 10 && sp->bit_depth != 12 && sp->bit_depth != 14)
        return fail(OHCU_EINVAL, "unsupported bit depth %d (8, 9, 10, 12, 14)", sp->bit_depth);
    if (sp->chroma_format_idc < 0 || sp->chroma_format_idc > 3)
        return fail(OHCU_EINVAL, "bad chroma_format_idc %d", sp->chroma_format_idc);
    if (sp->log2_ctb_size < 4 || sp->log2_ctb_size > 6 || sp->log2_min_cb_size < 3 ||
        sp->log2_min_cb_size > sp->log2_ctb_size || sp->log2_min_pu_size < 2 ||
        sp->log2_min_pu_size > sp->log2_min_cb_size)
        return fail(OHCU_EINVAL, "bad block size parameters");
    if (sp->log2_ctb_size == 4 && (sp->chroma_format_idc == 1 || sp->chroma_format_idc == 2))
        return fail(OHCU_EINVAL, "16x16 CTBs with subsampled chroma are not supported "
                                 "(reference SAO reads chroma before its horizontal edges are filtered)");
    if ((sp->width & ((1 << sp->log2_min_cb_size) - 1)) || (sp->height & ((1 << sp->log2_min_cb_size) - 1)))
        return fail(OHCU_EINVAL, "picture size must be a multiple of the minimum CB size");
    // ff_hevc_pred_init (hevcpred.c:62-76) has predictors for 8, 9, 10 and 12 bits only
    if (sp->intra && sp->bit_depth > 12)
        return fail(OHCU_EINVAL, "intra prediction at %d bits: the reference instantiates its predictors for 8, 9, 10 and 12 bits", sp->bit_depth);
    if (sp->dpb_slots < 1 || sp->dpb_slots > OHCU_MAX_DPB) return fail(OHCU_EINVAL, "dpb_slots out of range");
    if (sp->jobs < 1 || sp->jobs > 64) return fail(OHCU_EINVAL, "jobs out of range (1..64)");

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) return fail(OHCU_ENODEV, "no CUDA device");
    if (device < 0 || device >= ndev) return fail(OHCU_ENODEV, "device %d of %d", device, ndev);
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(OHCU_ENODEV, "device is sm_%d%d; this library is built for sm_100a only", prop.major, prop.minor);

    ohcu_ctx *c = new (std::nothrow) ohcu_ctx();
    if (!c) return fail(OHCU_ENOMEM, "out of memory");
    c->device = device;
    c->sp = *sp;
    c->hs = (sp->chroma_format_idc == 1 || sp->chroma_format_idc == 2);
    c->vs = (sp->chroma_format_idc == 1);
    c->nplanes = sp->chroma_format_idc ? 3 : 1;
    c->P = sp->bit_depth > 8 ? 2 : 1;
    for (int i = 0; i < 3; i++) {
        c->pw[i] = i ? sp->width >> c->hs : sp->width;
        c->ph[i] = i ? sp->height >> c->vs : sp->height;
    }
    const int ctb = 1 << sp->log2_ctb_size;
    c->ctb_w = (sp->width + ctb - 1) >> sp->log2_ctb_size;
    c->ctb_h = (sp->height + ctb - 1) >> sp->log2_ctb_size;
    c->nctb = c->ctb_w * c->ctb_h;
    c->min_cb_w = sp->width >> sp->log2_min_cb_size;  c->min_cb_h = sp->height >> sp->log2_min_cb_size;
    c->min_pu_w = sp->width >> sp->log2_min_pu_size;  c->min_pu_h = sp->height >> sp->log2_min_pu_size;
    c->bs_w = (sp->width >> 2) + 1;                   c->bs_h = (sp->height >> 2) + 1;

    // worst-case list sizes: two 8x4 PUs per 8x8 CU; one 4x4 TB per 16 samples; one int16 per sample
    const size_t nsamp = (size_t)c->pw[0] * c->ph[0] + (c->nplanes == 3 ? 2 * (size_t)c->pw[1] * c->ph[1] : 0);
    if (c->sp.max_pus <= 0) c->sp.max_pus = (int)(((size_t)sp->width * sp->height) / 32);
    if (c->sp.max_tus <= 0) c->sp.max_tus = (int)(nsamp / 16);
    if (c->sp.max_coeffs <= 0) c->sp.max_coeffs = (int64_t)nsamp;
    if (c->sp.max_ius <= 0) c->sp.max_ius = (int)(nsamp / 16);             // one predicted block per 4x4

    int rc = 0;
#define TRY(x) do { rc = (x); if (rc) { ohcu_destroy(c); return rc; } } while (0)
#define TRYCU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { ohcu_destroy(c); \
        return fail(e_ == cudaErrorMemoryAllocation ? OHCU_ENOMEM : OHCU_ECUDA, "%s -> %s", #x, cudaGetErrorString(e_)); } } while (0)
    if (upload_dct_table() || upload_mc_taps()) { ohcu_destroy(c); return fail(OHCU_ECUDA, "constant upload failed"); }
    if (recon_set_attrs() || loopfilter_set_attrs() || intra_set_attrs()) { ohcu_destroy(c); return fail(OHCU_ECUDA, "cudaFuncSetAttribute failed"); }
    c->slots.resize(sp->dpb_slots);
    for (auto &s : c->slots) {
        TRY(alloc_picture(c, &s.pic, true));
        TRYCU(cudaEventCreateWithFlags(&s.ready, cudaEventDisableTiming));
        TRYCU(cudaEventCreateWithFlags(&s.read_done, cudaEventDisableTiming));
    }
    c->jobs.resize(sp->jobs);
    for (auto &j : c->jobs) {
        TRYCU(cudaStreamCreateWithFlags(&j.stream, cudaStreamNonBlocking));
        TRYCU(cudaEventCreateWithFlags(&j.done, cudaEventDisableTiming));
        TRYCU(cudaStreamCreateWithFlags(&j.side, cudaStreamNonBlocking));
        TRYCU(cudaEventCreateWithFlags(&j.fork, cudaEventDisableTiming));
        TRYCU(cudaEventCreateWithFlags(&j.join, cudaEventDisableTiming));
        for (auto &e : j.ev) TRYCU(cudaEventCreate(&e));
        TRY(alloc_picture(c, &j.recon, false));
        TRY(alloc_arena(&j.pus, c->sp.max_pus));
        TRY(alloc_arena(&j.tus, c->sp.max_tus));
        TRY(alloc_arena(&j.tuq, c->sp.max_tus));
        TRYCU(cudaHostAlloc((void **)&j.sl_h, SL_BYTES, cudaHostAllocDefault));
        TRYCU(cudaMalloc((void **)&j.sl_d, SL_BYTES));
        TRYCU(cudaHostAlloc((void **)&j.err_h, sizeof(int), cudaHostAllocDefault));
        TRYCU(cudaMalloc((void **)&j.err_d, sizeof(int)));
        TRYCU(cudaMemset(j.err_d, 0, sizeof(int)));
        *j.err_h = 0;
        if (c->sp.intra) {
            TRY(alloc_arena(&j.ius, c->sp.max_ius));
            TRY(alloc_arena(&j.iu_ctbs, c->nctb));
            memset(j.iu_ctbs.h, 0, sizeof(ohcu_iu_index) * c->nctb);
            size_t off[3], total = 0;
            for (int i = 0; i < c->nplanes; i++) {
                j.resid.pitch[i] = (c->pw[i] * 2 + 255) & ~255;
                off[i] = total;
                total += (size_t)j.resid.pitch[i] * c->ph[i];
            }
            TRYCU(cudaMalloc((void **)&j.resid_base, total));
            TRYCU(cudaMemset(j.resid_base, 0, total));
            for (int i = 0; i < c->nplanes; i++) j.resid.p[i] = j.resid_base + off[i];
            TRYCU(cudaMalloc((void **)&j.progress_d, sizeof(int) * (c->nctb + 1)));     // ticket + one flag per CTB
        }
        // twice max_coeffs int16: the sparse form of a frame whose coefficients are all significant
        TRY(alloc_arena(&j.coeffs, 2 * (size_t)c->sp.max_coeffs));
        j.ctb_area.assign(c->nctb, 0);
        for (int pass = 0; pass < 2; pass++) {            // pass 0 sizes the table block, pass 1 carves it
            size_t off = 0;
            carve_arena(&j.ctb_range, c->nctb, j.tab_h, j.tab_d, &off);
            carve_arena(&j.ctb_tu_range, c->nctb, j.tab_h, j.tab_d, &off);
            carve_arena(&j.ctb_coef_range, c->nctb, j.tab_h, j.tab_d, &off);
            carve_arena(&j.slices, OHCU_MAX_SLICES, j.tab_h, j.tab_d, &off);
            carve_arena(&j.ctbs, c->nctb, j.tab_h, j.tab_d, &off);
            carve_arena(&j.fse, c->nctb, j.tab_h, j.tab_d, &off);
            carve_arena(&j.dbk_disabled, c->nctb, j.tab_h, j.tab_d, &off);
            carve_arena(&j.tsa, c->nctb, j.tab_h, j.tab_d, &off);
            carve_arena(&j.tile, c->nctb, j.tab_h, j.tab_d, &off);
            j.tab_end[0] = off;
            carve_arena(&j.sao, c->nctb, j.tab_h, j.tab_d, &off);
            carve_arena(&j.sao_flags, c->nctb, j.tab_h, j.tab_d, &off);
            j.tab_end[1] = off;
            carve_arena(&j.qp, (size_t)c->min_cb_w * c->min_cb_h, j.tab_h, j.tab_d, &off);
            carve_arena(&j.db, c->nctb, j.tab_h, j.tab_d, &off);
            carve_arena(&j.bs_flags, c->nctb, j.tab_h, j.tab_d, &off);
            j.tab_end[2] = off;
            carve_arena(&j.vbs, (size_t)c->bs_w * c->bs_h, j.tab_h, j.tab_d, &off);
            carve_arena(&j.hbs, (size_t)c->bs_w * c->bs_h, j.tab_h, j.tab_d, &off);
            j.tab_end[3] = off;
            carve_arena(&j.is_pcm, (size_t)c->min_pu_w * c->min_pu_h, j.tab_h, j.tab_d, &off);
            j.tab_end[4] = off;
            if (pass == 0) {
                TRYCU(cudaHostAlloc((void **)&j.tab_h, off, cudaHostAllocDefault));
                TRYCU(cudaMalloc((void **)&j.tab_d, off));
                memset(j.tab_h, 0, off);
            }
        }
    }
    if (c->sp.intra) {
        // intra_kernel takes CTBs along anti-diagonals x + 2 y: every CTB a block can read (left, up-left, up, up-right)
        // lies on an earlier one, and a limited number of CTAs then covers whole wavefronts instead of a few CTB rows
        std::vector<int> order;
        order.reserve(c->nctb);
        for (int d = 0; d <= c->ctb_w - 1 + 2 * (c->ctb_h - 1); d++)
            for (int y = std::max(0, (d - (c->ctb_w - 1) + 1) / 2); y <= std::min(c->ctb_h - 1, d / 2); y++)
                order.push_back(y * c->ctb_w + d - 2 * y);
        if ((int)order.size() != c->nctb) { ohcu_destroy(c); return fail(OHCU_ECUDA, "internal: wavefront order"); }
        TRYCU(cudaMalloc((void **)&c->intra_order_d, sizeof(int) * c->nctb));
        TRYCU(cudaMemcpy(c->intra_order_d, order.data(), sizeof(int) * c->nctb, cudaMemcpyHostToDevice));
    }
    TRYCU(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    c->copy_lanes[0] = c->copy_stream;
    for (int i = 1; i < COPY_LANES; i++) TRYCU(cudaStreamCreateWithFlags(&c->copy_lanes[i], cudaStreamNonBlocking));
    for (size_t i = 0; i < c->slots.size(); i++) c->slots[i].copy_lane = (int)(i % COPY_LANES);
    TRY(build_tensor_maps(c));
#undef TRY
#undef TRYCU
    *out = c;
    return 0;
}

extern "C" int ohcu_get_seq_params(const ohcu_ctx *c, ohcu_seq_params *out)
{
    if (!c || !out) return fail(OHCU_EINVAL, "null argument");
    *out = c->sp;
    return 0;
}

// sequence-level part of the kernel parameter block
static void fill_seq_dev(const ohcu_ctx *c, FrameDev *f)
{
    memset(f, 0, sizeof(*f));
    const ohcu_seq_params &sp = c->sp;
    f->width = sp.width; f->height = sp.height; f->bd = sp.bit_depth; f->cfi = sp.chroma_format_idc;
    f->hs = c->hs; f->vs = c->vs;
    f->log2_ctb = sp.log2_ctb_size; f->ctb_w = c->ctb_w; f->ctb_h = c->ctb_h;
    f->log2_min_cb = sp.log2_min_cb_size; f->min_cb_w = c->min_cb_w;
    f->log2_min_pu = sp.log2_min_pu_size; f->min_pu_w = c->min_pu_w; f->min_pu_h = c->min_pu_h;
    f->bs_width = c->bs_w;
    f->pcm_or_bypass = sp.pcm_or_bypass; f->cb_qp_offset = sp.cb_qp_offset; f->cr_qp_offset = sp.cr_qp_offset;
    f->nplanes = c->nplanes;
    for (int i = 0; i < 3; i++) { f->pw[i] = c->pw[i]; f->ph[i] = c->ph[i]; }
    for (size_t i = 0; i < c->slots.size(); i++) f->dpb[i] = c->slots[i].pic.ps;
}

// ------------------------------------------------------------------ device DPB
static int check_slot(ohcu_ctx *c, int slot)
{
    if (slot < 0 || slot >= (int)c->slots.size()) return fail(OHCU_EINVAL, "bad DPB slot %d", slot);
    return 0;
}

// The error word prepare_kernel left for a job (valid once the job's stream has drained).
static int job_list_error(const ohcu_ctx *c, int job)
{
    const int e = *c->jobs[job].err_h;
    if (!e) return 0;
    return fail(OHCU_EINVAL, "job %d: %s %d); the frame was not reconstructed", job, prepare_error_text(e), (int)((unsigned)e >> 8));
}
static int slot_list_error(ohcu_ctx *c, int slot)
{
    int prod;
    {
        std::lock_guard<std::mutex> lock(c->mu);
        prod = c->slots[slot].producer;
        if (prod >= 0 && c->jobs[prod].out_slot != slot) prod = -1;      // the job moved on to another picture
    }
    return prod >= 0 ? job_list_error(c, prod) : 0;
}

extern "C" int ohcu_slot_upload(ohcu_ctx *c, int slot, const uint8_t *const planes[3], const ptrdiff_t strides[3])
{
    ENTER(c);
    if (check_slot(c, slot)) return OHCU_EINVAL;
    if (!planes || !strides) return fail(OHCU_EINVAL, "null planes");
    Slot &s = c->slots[slot];
    std::lock_guard<std::mutex> lock(c->mu);
    // anything still reading or writing this slot must finish first
    CU(cudaDeviceSynchronize());
    for (int i = 0; i < c->nplanes; i++)
        CU(cudaMemcpy2D(s.pic.ps.p[i], s.pic.ps.pitch[i], planes[i], strides[i], (size_t)c->pw[i] * c->P,
                        c->ph[i], cudaMemcpyHostToDevice));
    // replicated halo around the uploaded picture (MC windows are fetched from it)
    FrameDev f;
    fill_seq_dev(c, &f);
    launch_pad_picture(f, s.pic.ps, 0);
    CU(cudaGetLastError());
    CU(cudaDeviceSynchronize());
    c->launches++;
    s.has_ready = false;
    s.readers = 0;
    s.producer = -1;
    s.released = false;
    return 0;
}

extern "C" int ohcu_slot_pitch(const ohcu_ctx *c, int plane, ptrdiff_t *pitch)
{
    if (!c || !pitch || plane < 0 || plane >= c->nplanes) return fail(OHCU_EINVAL, "bad plane");
    *pitch = c->slots[0].pic.ps.pitch[plane];
    return 0;
}

static int download_impl(ohcu_ctx *c, int slot, uint8_t *const planes[3], const ptrdiff_t strides[3], bool wait)
{
    ENTER(c);
    if (check_slot(c, slot)) return OHCU_EINVAL;
    if (!planes || !strides) return fail(OHCU_EINVAL, "null planes");
    Slot &s = c->slots[slot];
    {
        std::lock_guard<std::mutex> lock(c->mu);
        if (s.released) return fail(OHCU_ESTATE, "DPB slot %d was released and not written since", slot);
        cudaStream_t cs = c->copy_lanes[s.copy_lane];       // downloads of one slot stay ordered on its lane
        if (s.has_ready) CU(cudaStreamWaitEvent(cs, s.ready, 0));
        for (int i = 0; i < c->nplanes; i++) {
            // a host plane with the slot's own pitch (ohcu_slot_pitch) is one linear copy: the row
            // tails carry halo bytes, and the copy engine runs at full rate
            if (strides[i] == (ptrdiff_t)s.pic.ps.pitch[i])
                CU(cudaMemcpyAsync(planes[i], s.pic.ps.p[i], (size_t)(c->ph[i] - 1) * s.pic.ps.pitch[i] + (size_t)c->pw[i] * c->P,
                                   cudaMemcpyDeviceToHost, cs));
            else
                CU(cudaMemcpy2DAsync(planes[i], strides[i], s.pic.ps.p[i], s.pic.ps.pitch[i], (size_t)c->pw[i] * c->P,
                                     c->ph[i], cudaMemcpyDeviceToHost, cs));
        }
        CU(cudaEventRecord(s.read_done, cs));
        s.has_read = true;
    }
    if (wait) CU(cudaEventSynchronize(s.read_done));
    return 0;
}
extern "C" int ohcu_slot_download(ohcu_ctx *c, int slot, uint8_t *const planes[3], const ptrdiff_t strides[3])
{
    return download_impl(c, slot, planes, strides, true);
}
extern "C" int ohcu_slot_download_async(ohcu_ctx *c, int slot, uint8_t *const planes[3], const ptrdiff_t strides[3])
{
    return download_impl(c, slot, planes, strides, false);
}
// Cropped copy-out: pack kernel into the slot's tight staging buffer, then linear copies.
static int output_cropped_impl(ohcu_ctx *c, int slot, const ohcu_crop *crop, uint8_t *const planes[3], bool wait)
{
    ENTER(c);
    if (check_slot(c, slot)) return OHCU_EINVAL;
    if (!crop || !planes) return fail(OHCU_EINVAL, "null crop or planes");
    const int mx = (1 << c->hs) - 1, my = (1 << c->vs) - 1;
    if (crop->left < 0 || crop->top < 0 || crop->width <= 0 || crop->height <= 0 ||
        crop->left + crop->width > c->sp.width || crop->top + crop->height > c->sp.height)
        return fail(OHCU_EINVAL, "crop window outside the picture");
    if (c->nplanes > 1 && (((crop->left | crop->width) & mx) || ((crop->top | crop->height) & my)))
        return fail(OHCU_EINVAL, "crop window not a multiple of the chroma subsampling");
    int left[3], top[3], w[3], h[3];
    size_t off[3], total = 0, full = 0;
    for (int i = 0; i < c->nplanes; i++) {
        if (!planes[i]) return fail(OHCU_EINVAL, "null plane");
        const int hs = i ? c->hs : 0, vs = i ? c->vs : 0;
        left[i] = crop->left >> hs; top[i] = crop->top >> vs; w[i] = crop->width >> hs; h[i] = crop->height >> vs;
        off[i] = total;
        total += ((size_t)w[i] * h[i] * c->P + 15) & ~(size_t)15;     // planes start on 16-byte boundaries
        full += ((size_t)c->pw[i] * c->ph[i] * c->P + 15) & ~(size_t)15;
    }
    Slot &s = c->slots[slot];
    {
        std::lock_guard<std::mutex> lock(c->mu);
        if (s.released) return fail(OHCU_ESTATE, "DPB slot %d was released and not written since", slot);
        cudaStream_t cs = c->copy_lanes[s.copy_lane];
        if (!s.out_d) CU(cudaMalloc(&s.out_d, full));
        if (s.has_ready) CU(cudaStreamWaitEvent(cs, s.ready, 0));
        uint8_t *dst[3] = {nullptr, nullptr, nullptr};
        for (int i = 0; i < c->nplanes; i++) dst[i] = s.out_d + off[i];
        c->launches += launch_crop_pack(s.pic.ps, c->nplanes, c->P, left, top, w, h, dst, cs);
        CU(cudaGetLastError());
        // planes that follow each other in the caller's memory the way they do in the staging
        // buffer leave as one copy
        bool one = true;
        for (int i = 1; i < c->nplanes; i++) one = one && planes[i] == planes[0] + off[i];
        if (one) {
            CU(cudaMemcpyAsync(planes[0], s.out_d, off[c->nplanes - 1] + (size_t)w[c->nplanes - 1] * h[c->nplanes - 1] * c->P,
                               cudaMemcpyDeviceToHost, cs));
        } else {
            for (int i = 0; i < c->nplanes; i++)
                CU(cudaMemcpyAsync(planes[i], dst[i], (size_t)w[i] * h[i] * c->P, cudaMemcpyDeviceToHost, cs));
        }
        CU(cudaEventRecord(s.read_done, cs));
        s.has_read = true;
    }
    if (wait) CU(cudaEventSynchronize(s.read_done));
    return 0;
}
extern "C" int ohcu_slot_output_cropped(ohcu_ctx *c, int slot, const ohcu_crop *crop, uint8_t *const planes[3])
{
    return output_cropped_impl(c, slot, crop, planes, true);
}
extern "C" int ohcu_slot_output_cropped_async(ohcu_ctx *c, int slot, const ohcu_crop *crop, uint8_t *const planes[3])
{
    return output_cropped_impl(c, slot, crop, planes, false);
}
// verify_md5 (hevcdec.c:4035-4098) without moving the picture: one MD5 per plane, computed on the device
extern "C" int ohcu_slot_md5(ohcu_ctx *c, int slot, uint8_t md5[3][16])
{
    ENTER(c);
    if (check_slot(c, slot)) return OHCU_EINVAL;
    if (!md5) return fail(OHCU_EINVAL, "null argument");
    Slot &s = c->slots[slot];
    uint8_t *out = nullptr;
    cudaStream_t cs;
    {
        std::lock_guard<std::mutex> lock(c->mu);
        if (s.released) return fail(OHCU_ESTATE, "DPB slot %d was released and not written since", slot);
        if (!c->md5_d) CU(cudaMalloc((void **)&c->md5_d, 48 * c->slots.size()));
        out = c->md5_d + 48 * slot;
        cs = c->copy_lanes[s.copy_lane];
        if (s.has_ready) CU(cudaStreamWaitEvent(cs, s.ready, 0));
        c->launches += launch_md5(s.pic.ps, c->nplanes, c->P, c->pw, c->ph, out, cs);
        CU(cudaGetLastError());
        CU(cudaEventRecord(s.read_done, cs));                  // the next writer of the slot waits for the hash like for a download
        s.has_read = true;
    }
    memset(md5, 0, 48);
    CU(cudaMemcpyAsync(md5, out, 16 * c->nplanes, cudaMemcpyDeviceToHost, cs));
    CU(cudaStreamSynchronize(cs));
    return slot_list_error(c, slot);
}

extern "C" int ohcu_slot_sync(ohcu_ctx *c, int slot)
{
    ENTER(c);
    if (check_slot(c, slot)) return OHCU_EINVAL;
    Slot &s = c->slots[slot];
    bool ready, read;
    {
        std::lock_guard<std::mutex> lock(c->mu);
        ready = s.has_ready; read = s.has_read;
    }
    if (ready) CU(cudaEventSynchronize(s.ready));
    if (read) CU(cudaEventSynchronize(s.read_done));
    return slot_list_error(c, slot);
}

// hevc_refs.c:131-166: the decoder is done with the picture in `slot`
extern "C" int ohcu_slot_release(ohcu_ctx *c, int slot)
{
    ENTER(c);
    if (check_slot(c, slot)) return OHCU_EINVAL;
    std::lock_guard<std::mutex> lock(c->mu);
    c->slots[slot].released = true;
    return 0;
}

// ------------------------------------------------------------------ per-frame lists
extern "C" int ohcu_frame_begin(ohcu_ctx *c, int out_slot, int poc)
{
    ENTER(c);
    if (check_slot(c, out_slot)) return OHCU_EINVAL;
    int id;
    {
        std::lock_guard<std::mutex> lock(c->mu);
        id = c->next_job;
        if (c->jobs[id].open) return fail(OHCU_ESTATE, "job %d is still being filled: more frames begun than jobs", id);
        c->next_job = (c->next_job + 1) % (int)c->jobs.size();
        c->jobs[id].open = true;
    }
    Job &j = c->jobs[id];
    // the arena's previous frame must have left the device before its pinned staging is reused
    if (j.has_done) CU(cudaEventSynchronize(j.done));
    j.open = true; j.staged = false;
    j.out_slot = out_slot; j.poc = poc;
    j.n_pu = 0; j.n_tu = 0; j.ncoeff = 0; j.n_slices = 0;
    j.pu_sorted = true; j.last_ctb = -1;
    j.tu_sorted = true; j.last_tu_ctb = -1;
    j.has_cross = false; j.max_slice = -1;
    j.has_deblock = j.has_sao = j.has_pcm = false;
    memset(j.ctb_range.h, 0, sizeof(int2) * c->nctb);
    memset(j.ctb_tu_range.h, 0, sizeof(int2) * c->nctb);
    j.coef_form = 0; j.has_sl = false;
    std::fill(j.ctb_area.begin(), j.ctb_area.end(), 0);
    j.pp.pcm_or_bypass = c->sp.pcm_or_bypass; j.pp.cb_qp_offset = c->sp.cb_qp_offset;
    j.pp.cr_qp_offset = c->sp.cr_qp_offset; j.pp.no_tile_filter = c->sp.no_tile_filter;
    j.inplace = false; j.all_inter = 0; j.n_coef_units = 0; j.n_pcm = 0; j.pcm_bytes = 0;
    memset(j.ctbs.h, 0, sizeof(ohcu_ctb_index) * c->nctb);
    if (j.n_iu || j.inplace_iu) memset(j.iu_ctbs.h, 0, sizeof(ohcu_iu_index) * c->nctb);
    j.n_iu = 0; j.last_iu_ctb = -1; j.inplace_iu = false;
    *j.err_h = 0;
    return id;
}

extern "C" int ohcu_set_pic_params(ohcu_ctx *c, int job, const ohcu_pic_params *pp)
{
    ENTER(c);
    if (check_job(c, job)) return OHCU_EINVAL;
    Job &j = c->jobs[job];
    if (!j.open) return fail(OHCU_ESTATE, "job %d is not open (call ohcu_frame_begin)", job);
    if (!pp) return fail(OHCU_EINVAL, "null picture parameters");
    j.pp = *pp;
    return 0;
}

extern "C" int ohcu_frame_abort(ohcu_ctx *c, int job)
{
    ENTER(c);
    if (check_job(c, job)) return OHCU_EINVAL;
    c->jobs[job].open = false;
    return 0;
}

#define OPEN_JOB(c, job)                                                        \
    ENTER(c);                                                                   \
    if (check_job(c, job)) return OHCU_EINVAL;                                  \
    Job &j = (c)->jobs[job];                                                    \
    if (!j.open) return fail(OHCU_ESTATE, "job %d is not open (call ohcu_frame_begin)", job)

extern "C" int ohcu_set_slice(ohcu_ctx *c, int job, int slice_idx, const ohcu_slice_tab *tab)
{
    OPEN_JOB(c, job);
    if (!tab || slice_idx < 0 || slice_idx >= OHCU_MAX_SLICES) return fail(OHCU_EINVAL, "bad slice index %d", slice_idx);
    j.slices.h[slice_idx] = *tab;
    if (slice_idx + 1 > j.n_slices) j.n_slices = slice_idx + 1;
    return 0;
}

extern "C" int ohcu_append_pu(ohcu_ctx *c, int job, const ohcu_pu *recs, int n)
{
    OPEN_JOB(c, job);
    if (n < 0 || (n && !recs)) return fail(OHCU_EINVAL, "bad PU list");
    if ((size_t)j.n_pu + n > j.pus.cap) return fail(OHCU_EOVERFLOW, "PU arena full (%zu)", j.pus.cap);
    const int l2 = c->sp.log2_ctb_size, ctb_w = c->ctb_w, W = c->sp.width, H = c->sp.height, base = j.n_pu;
    int2 *const range = j.ctb_range.h;
    int *const area = j.ctb_area.data();
    ohcu_pu *const dst = j.pus.h + j.n_pu;
    int last = j.last_ctb, max_slice = j.max_slice;
    bool sorted = j.pu_sorted;
    for (int i = 0; i < n; i++) {
        const ohcu_pu pu = recs[i];
        // a block that is not 4-aligned or crosses a CTB boundary would leave the CTB's 16x16 cell map (mc_kernel)
        if (pu.pred_flag < 1 || pu.pred_flag > 3 || pu.w < 4 || pu.h < 4 || pu.w > 64 || pu.h > 64 ||
            ((pu.x0 | pu.y0 | pu.w | pu.h) & 3) || pu.x0 + pu.w > W || pu.y0 + pu.h > H ||
            (pu.x0 >> l2) != ((pu.x0 + pu.w - 1) >> l2) || (pu.y0 >> l2) != ((pu.y0 + pu.h - 1) >> l2) ||
            pu.slice >= OHCU_MAX_SLICES || pu.ref_idx[0] >= OHCU_MAX_REFS || pu.ref_idx[1] >= OHCU_MAX_REFS)
            return fail(OHCU_EINVAL, "PU record %d is malformed", base + i);
        const int ctb = (pu.x0 >> l2) + (pu.y0 >> l2) * ctb_w;
        if (ctb != last) {
            if (range[ctb].y != 0) sorted = false;           // CTB seen before: lists not grouped
            else range[ctb].x = base + i;
            last = ctb;
        }
        range[ctb].y++;
        area[ctb] += pu.w * pu.h;
        max_slice = std::max(max_slice, (int)pu.slice);
        dst[i] = pu;
    }
    j.last_ctb = last; j.pu_sorted = sorted; j.max_slice = max_slice;
    j.n_pu += n;
    return 0;
}

extern "C" int ohcu_append_iu(ohcu_ctx *c, int job, const ohcu_iu *recs, int n)
{
    OPEN_JOB(c, job);
    if (n < 0 || (n && !recs)) return fail(OHCU_EINVAL, "bad intra block list");
    if (!c->sp.intra) return fail(OHCU_ESTATE, "the context was created without ohcu_seq_params.intra");
    if ((size_t)j.n_iu + n > j.ius.cap) return fail(OHCU_EOVERFLOW, "intra block arena full (%zu)", j.ius.cap);
    const int l2 = c->sp.log2_ctb_size, base = j.n_iu;
    ohcu_iu_index *const idx = j.iu_ctbs.h;
    int last = j.last_iu_ctb;
    for (int i = 0; i < n; i++) {
        const ohcu_iu u = recs[i];
        if (!iu_ok(u, c->nplanes, c->pw, c->ph, c->hs, c->vs, l2))
            return fail(OHCU_EINVAL, "intra block record %d is malformed (size, mode, position, or a neighbour named available that cannot be)", base + i);
        const int hs = u.c_idx ? c->hs : 0, vs = u.c_idx ? c->vs : 0;
        const int ctb = ((u.x0 << hs) >> l2) + ((u.y0 << vs) >> l2) * c->ctb_w;
        if (ctb != last) {
            if (idx[ctb].n_iu) return fail(OHCU_EINVAL, "intra block record %d: the blocks of a CTB must be appended contiguously", base + i);
            idx[ctb].first_iu = (uint32_t)(base + i);
            last = ctb;
        }
        if (idx[ctb].n_iu >= 768u) return fail(OHCU_EINVAL, "intra block record %d: more than 768 blocks in CTB %d", base + i, ctb);
        idx[ctb].n_iu++;
        j.ius.h[base + i] = u;
    }
    j.last_iu_ctb = last;
    j.n_iu += n;
    return 0;
}

static int add_tus(ohcu_ctx *c, Job &j, const ohcu_tu *recs, int n, uint32_t rebase)
{
    if ((size_t)j.n_tu + n > j.tus.cap) return fail(OHCU_EOVERFLOW, "TU arena full (%zu)", j.tus.cap);
    // everything the loop reads, in locals: the stores into the pinned arena could alias the context
    int2 *const range = j.ctb_tu_range.h;
    ohcu_tu *const dst = j.tus.h + j.n_tu;
    const int l2 = c->sp.log2_ctb_size, ctb_w = c->ctb_w, base = j.n_tu;
    const int pw[3] = {c->pw[0], c->pw[1], c->pw[2]}, ph[3] = {c->ph[0], c->ph[1], c->ph[2]};
    const int ths[3] = {0, c->hs, c->hs}, tvs[3] = {0, c->vs, c->vs};
    const int max_c = c->sp.chroma_format_idc ? 2 : 0;
    const size_t ncoeff = j.ncoeff;
    const bool intra_ok = c->sp.intra != 0;          // OHCU_TUF_INTRA blocks need the job's residual planes
    int last = j.last_tu_ctb;
    bool sorted = j.tu_sorted, cross = j.has_cross;
    for (int i = 0; i < n; i++) {
        ohcu_tu tu = recs[i];
        const unsigned l = (unsigned)tu.log2_size - 2u;
        if (l > 3u || tu.c_idx > max_c || tu.mode > OHCU_TU_BYPASS || (tu.mode == OHCU_TU_DST4 && l != 0) ||
            ((tu.flags & OHCU_TUF_INTRA) && !intra_ok))
            return fail(OHCU_EINVAL, "TU record %d is malformed", base + i);
        const int n1 = 1 << tu.log2_size;
        if (tu.x0 + n1 > pw[tu.c_idx] || tu.y0 + n1 > ph[tu.c_idx] || ((tu.x0 | tu.y0) & 3) ||
            ((tu.x0 << ths[tu.c_idx]) >> l2) != (((tu.x0 + n1 - 1) << ths[tu.c_idx]) >> l2) ||
            ((tu.y0 << tvs[tu.c_idx]) >> l2) != (((tu.y0 + n1 - 1) << tvs[tu.c_idx]) >> l2))
            return fail(OHCU_EINVAL, "TU record %d lies outside plane %d, is not 4-aligned or crosses a CTB boundary", base + i, tu.c_idx);
        tu.coeff_off += rebase;
        if ((size_t)tu.coeff_off + (size_t)(n1 * n1) > ncoeff || (tu.coeff_off & 1))
            return fail(OHCU_EINVAL, "TU record %d: coefficient block outside the arena", base + i);
        // CTB of the block (plane samples -> luma samples)
        const int ctb = ((tu.x0 << ths[tu.c_idx]) >> l2) + ((tu.y0 << tvs[tu.c_idx]) >> l2) * ctb_w;
        if (ctb != last) {
            if (range[ctb].y != 0) sorted = false;           // CTB seen before: list not grouped
            else range[ctb].x = base + i;
            last = ctb;
        }
        range[ctb].y++;
        cross |= (tu.flags & OHCU_TUF_CROSS) != 0;
        dst[i] = tu;
    }
    j.last_tu_ctb = last; j.tu_sorted = sorted; j.has_cross = cross;
    j.n_tu += n;
    return 0;
}

extern "C" int ohcu_append_tu(ohcu_ctx *c, int job, const ohcu_tu *recs, int n, const int16_t *coeffs, size_t ncoeff)
{
    OPEN_JOB(c, job);
    if (n < 0 || (n && (!recs || !coeffs))) return fail(OHCU_EINVAL, "bad TU list");
    if (j.coef_form >= 2) return fail(OHCU_ESTATE, "job %d already holds sparse coefficients", job);
    j.coef_form = 1;
    if (j.ncoeff + ncoeff > j.coeffs.cap) return fail(OHCU_EOVERFLOW, "coefficient arena full (%zu)", j.coeffs.cap);
    if (j.ncoeff + ncoeff > 0xffffffffull) return fail(OHCU_EOVERFLOW, "coefficient offsets exceed 32 bits");
    const uint32_t base = (uint32_t)j.ncoeff;
    memcpy(j.coeffs.h + j.ncoeff, coeffs, ncoeff * sizeof(int16_t));
    j.ncoeff += ncoeff;
    return add_tus(c, j, recs, n, base);
}

extern "C" int ohcu_reserve_coeffs(ohcu_ctx *c, int job, size_t ncoeff, int16_t **ptr, uint32_t *arena_off)
{
    OPEN_JOB(c, job);
    if (!ptr || !arena_off) return fail(OHCU_EINVAL, "null argument");
    if (j.coef_form >= 2) return fail(OHCU_ESTATE, "job %d already holds sparse coefficients", job);
    j.coef_form = 1;
    ncoeff = (ncoeff + 7) & ~(size_t)7;              // keep blocks 16-byte aligned
    if (j.ncoeff + ncoeff > j.coeffs.cap) return fail(OHCU_EOVERFLOW, "coefficient arena full (%zu)", j.coeffs.cap);
    *ptr = j.coeffs.h + j.ncoeff;
    *arena_off = (uint32_t)j.ncoeff;
    j.ncoeff += ncoeff;
    return 0;
}

extern "C" int ohcu_append_tu_inplace(ohcu_ctx *c, int job, const ohcu_tu *recs, int n)
{
    OPEN_JOB(c, job);
    if (n < 0 || (n && !recs)) return fail(OHCU_EINVAL, "bad TU list");
    if (j.coef_form >= 2) return fail(OHCU_ESTATE, "job %d already holds sparse coefficients", job);
    j.coef_form = 1;
    return add_tus(c, j, recs, n, 0);
}

// ---- sparse form: (position, level) pairs, dense blocks rebuilt per CTB in shared memory
extern "C" int ohcu_reserve_coefs(ohcu_ctx *c, int job, size_t n, ohcu_coef **ptr)
{
    OPEN_JOB(c, job);
    if (!ptr) return fail(OHCU_EINVAL, "null argument");
    if (j.coef_form == 1) return fail(OHCU_ESTATE, "job %d already holds dense coefficients", job);
    if (j.ncoeff + 2 * n > j.coeffs.cap) return fail(OHCU_EOVERFLOW, "coefficient arena full (%zu pairs)", j.coeffs.cap / 2);
    *ptr = (ohcu_coef *)(j.coeffs.h + j.ncoeff);
    return 0;
}

// form: 2 = the pairs are coefficients, 3 = un-scaled levels with one ohcu_tu_quant per record
static int append_pairs(ohcu_ctx *c, int job, const ohcu_tu *recs, const ohcu_tu_quant *quant, int n, const ohcu_coef *coefs,
                        size_t ncoefs, int form)
{
    OPEN_JOB(c, job);
    if (n < 0 || (n && !recs) || (ncoefs && !coefs) || (form == 3 && n && !quant)) return fail(OHCU_EINVAL, "bad TU list");
    if (j.coef_form == 1) return fail(OHCU_ESTATE, "job %d already holds dense coefficients", job);
    if (j.coef_form && j.coef_form != form)
        return fail(OHCU_ESTATE, "job %d already holds %s", job, j.coef_form == 3 ? "un-scaled levels" : "scaled coefficients");
    for (int i = 0; form == 3 && i < n; i++)
        if (quant[i].qp >= 76 || (quant[i].matrix_id > 5 && quant[i].matrix_id != OHCU_QM_FLAT))
            return fail(OHCU_EINVAL, "quantiser of TU record %d is malformed", j.n_tu + i);
    if ((size_t)j.n_tu + n > j.tus.cap) return fail(OHCU_EOVERFLOW, "TU arena full (%zu)", j.tus.cap);
    if (j.ncoeff + 2 * ncoefs > j.coeffs.cap) return fail(OHCU_EOVERFLOW, "coefficient arena full (%zu pairs)", j.coeffs.cap / 2);
    int2 *const range = j.ctb_tu_range.h, *const crange = j.ctb_coef_range.h;
    ohcu_tu *const dst = j.tus.h + j.n_tu;
    const int l2 = c->sp.log2_ctb_size, ctb_w = c->ctb_w, base = j.n_tu;
    const int pw[3] = {c->pw[0], c->pw[1], c->pw[2]}, ph[3] = {c->ph[0], c->ph[1], c->ph[2]};
    const int ths[3] = {0, c->hs, c->hs}, tvs[3] = {0, c->vs, c->vs};
    const int max_c = c->sp.chroma_format_idc ? 2 : 0;
    const uint32_t area = (uint32_t)recon_sparse_ctb_coefs(c->sp.log2_ctb_size, c->nplanes, c->hs, c->vs);
    const size_t first_pair = j.ncoeff / 2;
    const bool intra_ok = c->sp.intra != 0;
    size_t cursor = 0;
    int last = j.last_tu_ctb;
    bool cross = j.has_cross;
    // validate everything before the lists change
    if (j.ctb_seen.size() != (size_t)c->nctb) j.ctb_seen.assign(c->nctb, 0);
    j.seen_epoch++;
    for (int i = 0; i < n; i++) {
        const ohcu_tu &tu = recs[i];
        const unsigned l = (unsigned)tu.log2_size - 2u;
        if (l > 3u || tu.c_idx > max_c || tu.mode > OHCU_TU_BYPASS || (tu.mode == OHCU_TU_DST4 && l != 0) ||
            ((tu.flags & OHCU_TUF_INTRA) && !intra_ok))
            return fail(OHCU_EINVAL, "TU record %d is malformed", base + i);
        const int n1 = 1 << tu.log2_size;
        if (tu.x0 + n1 > pw[tu.c_idx] || tu.y0 + n1 > ph[tu.c_idx] || ((tu.x0 | tu.y0) & 3) ||
            ((tu.x0 << ths[tu.c_idx]) >> l2) != (((tu.x0 + n1 - 1) << ths[tu.c_idx]) >> l2) ||
            ((tu.y0 << tvs[tu.c_idx]) >> l2) != (((tu.y0 + n1 - 1) << tvs[tu.c_idx]) >> l2))
            return fail(OHCU_EINVAL, "TU record %d lies outside plane %d, is not 4-aligned or crosses a CTB boundary", base + i, tu.c_idx);
        if ((tu.coeff_off & 7) || tu.coeff_off + (uint32_t)(n1 * n1) > area || tu.pad > n1 * n1)
            return fail(OHCU_EINVAL, "TU record %d: block outside the CTB's coefficient area (%u)", base + i, area);
        const int ctb = ((tu.x0 << ths[tu.c_idx]) >> l2) + ((tu.y0 << tvs[tu.c_idx]) >> l2) * ctb_w;
        if (ctb != last) {
            if (range[ctb].y != 0 || j.ctb_seen[ctb] == j.seen_epoch)
                return fail(OHCU_EINVAL, "TU record %d: the sparse form needs the blocks of a CTB appended contiguously", base + i);
            j.ctb_seen[ctb] = j.seen_epoch;
        }
        last = ctb;
        cursor += tu.pad;
    }
    if (cursor != ncoefs) return fail(OHCU_EINVAL, "the records count %zu coefficients, %zu given", cursor, ncoefs);
    ohcu_coef *const arena = (ohcu_coef *)(j.coeffs.h + j.ncoeff);
    if (ncoefs && coefs != arena) memcpy(arena, coefs, ncoefs * sizeof(ohcu_coef));
    cursor = first_pair;
    last = j.last_tu_ctb;
    for (int i = 0; i < n; i++) {
        const ohcu_tu &tu = recs[i];
        const int ctb = ((tu.x0 << ths[tu.c_idx]) >> l2) + ((tu.y0 << tvs[tu.c_idx]) >> l2) * ctb_w;
        if (ctb != last) { range[ctb].x = base + i; crange[ctb].x = (int)cursor; crange[ctb].y = 0; last = ctb; }
        range[ctb].y++;
        crange[ctb].y += tu.pad;
        cursor += tu.pad;
        cross |= (tu.flags & OHCU_TUF_CROSS) != 0;
        dst[i] = tu;
    }
    if (form == 3 && n) memcpy(j.tuq.h + j.n_tu, quant, (size_t)n * sizeof(ohcu_tu_quant));
    j.last_tu_ctb = last; j.has_cross = cross; j.coef_form = form;
    j.ncoeff += 2 * ncoefs;
    j.n_tu += n;
    return 0;
}

extern "C" int ohcu_append_tu_sparse(ohcu_ctx *c, int job, const ohcu_tu *recs, int n, const ohcu_coef *coefs, size_t ncoefs)
{
    return append_pairs(c, job, recs, nullptr, n, coefs, ncoefs, 2);
}

// scale_and_clip_coeff (hevc_cabac.c:2030-2055) moves to the device: see recon.cu, "levels"
extern "C" int ohcu_append_tu_levels(ohcu_ctx *c, int job, const ohcu_tu *recs, const ohcu_tu_quant *quant, int n,
                                     const ohcu_coef *levels, size_t nlevels)
{
    return append_pairs(c, job, recs, quant, n, levels, nlevels, 3);
}

extern "C" int ohcu_set_scaling_list(ohcu_ctx *c, int job, const uint8_t sl[4][6][64], const uint8_t sl_dc[2][6])
{
    OPEN_JOB(c, job);
    if (!sl || !sl_dc) return fail(OHCU_EINVAL, "null scaling list");
    memcpy(j.sl_h, sl, 4 * 6 * 64);
    memcpy(j.sl_h + 4 * 6 * 64, sl_dc, 2 * 6);
    j.has_sl = true;
    return 0;
}

extern "C" int ohcu_set_deblock(ohcu_ctx *c, int job, const ohcu_deblock_in *in)
{
    OPEN_JOB(c, job);
    if (!in || !in->vertical_bs || !in->horizontal_bs || !in->qp_y_tab || !in->deblock)
        return fail(OHCU_EINVAL, "null deblock input");
    memcpy(j.vbs.h, in->vertical_bs, (size_t)c->bs_w * c->bs_h);
    memcpy(j.hbs.h, in->horizontal_bs, (size_t)c->bs_w * c->bs_h);
    memcpy(j.qp.h, in->qp_y_tab, (size_t)c->min_cb_w * c->min_cb_h);
    memcpy(j.db.h, in->deblock, sizeof(ohcu_dbparams) * c->nctb);
    j.has_pcm = j.pp.pcm_or_bypass && in->is_pcm;
    if (j.has_pcm) memcpy(j.is_pcm.h, in->is_pcm, (size_t)c->min_pu_w * c->min_pu_h);
    j.has_deblock = true;
    j.bs_derived = false;
    return 0;
}

// ff_hevc_deblocking_boundary_strengths moves to the device (bs.cu); here only the per-CTB
// conditions of lc->boundary_flags (hevcdec.c:2822-2837) are folded into one byte per CTB
extern "C" int ohcu_set_deblock_derived(ohcu_ctx *c, int job, const ohcu_deblock_derive_in *in)
{
    OPEN_JOB(c, job);
    if (!in || !in->qp_y_tab || !in->deblock) return fail(OHCU_EINVAL, "null deblock input");
    memcpy(j.qp.h, in->qp_y_tab, (size_t)c->min_cb_w * c->min_cb_h);
    memcpy(j.db.h, in->deblock, sizeof(ohcu_dbparams) * c->nctb);
    j.has_pcm = j.pp.pcm_or_bypass && in->is_pcm;
    if (j.has_pcm) memcpy(j.is_pcm.h, in->is_pcm, (size_t)c->min_pu_w * c->min_pu_h);
    const int cw = c->ctb_w;
    for (int rs = 0; rs < c->nctb; rs++) {
        const int x = rs % cw, y = rs / cw;
        const bool across = in->filter_slice_edges ? in->filter_slice_edges[rs] != 0 : true;
        auto differs = [&](int other) {
            return (in->tab_slice_address && in->tab_slice_address[rs] != in->tab_slice_address[other] && !across) ||
                   (j.pp.no_tile_filter && in->tile_id && in->tile_id[rs] != in->tile_id[other]);
        };
        unsigned fl = 0;
        if (x > 0 && differs(rs - 1)) fl |= BSF_NO_LEFT;
        if (y > 0 && differs(rs - cw)) fl |= BSF_NO_UP;
        if (in->deblock_disabled && in->deblock_disabled[rs]) fl |= BSF_DISABLED;
        j.bs_flags.h[rs] = (uint8_t)fl;
    }
    j.has_deblock = true;
    j.bs_derived = true;
    return 0;
}

// per-CTB restore flags of sao_filter_CTB (hevc_filter.c:277-325)
extern "C" int ohcu_set_sao(ohcu_ctx *c, int job, const ohcu_sao_in *in)
{
    OPEN_JOB(c, job);
    if (!in || !in->sao) return fail(OHCU_EINVAL, "null SAO input");
    memcpy(j.sao.h, in->sao, sizeof(ohcu_sao) * c->nctb);
    const int cw = c->ctb_w, chh = c->ctb_h;
    const bool no_tile = j.pp.no_tile_filter != 0;
    for (int y = 0; y < chh; y++)
        for (int x = 0; x < cw; x++) {
            const int rs = y * cw + x;
            const bool lfase = in->filter_slice_edges ? in->filter_slice_edges[rs] != 0 : true;
            unsigned fl = 0;
            if (no_tile || !lfase) {
                fl |= SAOF_RESTORE;
                auto slice_at = [&](int dx, int dy) { return in->tab_slice_address ? in->tab_slice_address[rs + dy * cw + dx] : 0; };
                auto tile_at = [&](int dx, int dy) { return in->tile_id ? in->tile_id[rs + dy * cw + dx] : 0; };
                auto dslice = [&](int dx, int dy) { return !lfase && slice_at(0, 0) != slice_at(dx, dy); };
                auto dtile = [&](int dx, int dy) { return no_tile && tile_at(0, 0) != tile_at(dx, dy); };
                const bool e0 = x == 0, e1 = y == 0, e2 = x == cw - 1, e3 = y == chh - 1;
                bool lt = false, rt = false, ut = false, bt = false;
                if (!e0) { lt = dtile(-1, 0); if (dslice(-1, 0) || lt) fl |= SAOF_VERT0; }
                if (!e2) { rt = dtile(1, 0);  if (dslice(1, 0) || rt)  fl |= SAOF_VERT1; }
                if (!e1) { ut = dtile(0, -1); if (dslice(0, -1) || ut) fl |= SAOF_HORIZ0; }
                if (!e3) { bt = dtile(0, 1);  if (dslice(0, 1) || bt)  fl |= SAOF_HORIZ1; }
                if (!e0 && !e1 && (dslice(-1, -1) || lt || ut)) fl |= SAOF_DIAG0;
                if (!e1 && !e2 && (dslice(1, -1) || rt || ut))  fl |= SAOF_DIAG1;
                if (!e2 && !e3 && (dslice(1, 1) || rt || bt))   fl |= SAOF_DIAG2;
                if (!e0 && !e3 && (dslice(-1, 1) || lt || bt))  fl |= SAOF_DIAG3;
            }
            j.sao_flags.h[rs] = (uint16_t)fl;
        }
    j.has_sao = true;
    return 0;
}

// ------------------------------------------------------------------ in-place lists
extern "C" int ohcu_frame_arenas_get(ohcu_ctx *c, int job, ohcu_frame_arenas *out)
{
    OPEN_JOB(c, job);
    if (!out) return fail(OHCU_EINVAL, "null argument");
    memset(out, 0, sizeof(*out));
    out->pus = j.pus.h;       out->max_pus = (uint32_t)j.pus.cap;
    out->tus = j.tus.h;       out->max_tus = (uint32_t)j.tus.cap;
    out->quant = j.tuq.h;
    out->coefs = (ohcu_coef *)j.coeffs.h;  out->coeffs = j.coeffs.h;  out->max_coefs = j.coeffs.cap / 2;
    out->ctbs = j.ctbs.h;     out->slices = j.slices.h;
    out->vertical_bs = j.vbs.h; out->horizontal_bs = j.hbs.h; out->qp_y_tab = j.qp.h; out->deblock = j.db.h;
    out->is_pcm = j.is_pcm.h; out->sao = j.sao.h; out->filter_slice_edges = j.fse.h;
    out->tab_slice_address = j.tsa.h; out->tile_id = j.tile.h; out->deblock_disabled = j.dbk_disabled.h;
    out->scaling_list = j.sl_h;
    if (c->sp.intra) {
        out->ius = j.ius.h; out->iu_ctbs = j.iu_ctbs.h; out->max_ius = (uint32_t)j.ius.cap;
        j.inplace_iu = true;
    }
    return 0;
}

extern "C" int ohcu_frame_commit_lists(ohcu_ctx *c, int job, const ohcu_frame_counts *n)
{
    OPEN_JOB(c, job);
    if (!n) return fail(OHCU_EINVAL, "null argument");
    if (j.n_pu || j.n_tu || j.n_iu) return fail(OHCU_ESTATE, "job %d already holds appended lists", job);
    if (n->n_iu && !c->sp.intra) return fail(OHCU_ESTATE, "the context was created without ohcu_seq_params.intra");
    if (n->n_iu > j.ius.cap) return fail(OHCU_EOVERFLOW, "intra block list larger than the job's arena");
    if (n->n_iu && n->all_inter) return fail(OHCU_EINVAL, "all_inter promised for a picture with intra blocks");
    if (n->n_pu > j.pus.cap || n->n_tu > j.tus.cap) return fail(OHCU_EOVERFLOW, "lists larger than the job's arenas");
    if (n->coef_form < 0 || n->coef_form > 3 || (n->n_tu && !n->coef_form)) return fail(OHCU_EINVAL, "bad coef_form %d", n->coef_form);
    const uint64_t n16 = n->coef_form >= 2 ? 2 * n->n_coef : n->n_coef;
    if (n16 > j.coeffs.cap || n16 > 0xffffffffull) return fail(OHCU_EOVERFLOW, "coefficient arena full (%zu int16)", j.coeffs.cap);
    if (n->n_slices < 0 || n->n_slices > OHCU_MAX_SLICES || (n->n_pu && !n->n_slices)) return fail(OHCU_EINVAL, "bad n_slices %d", n->n_slices);
    if (n->deblock < 0 || n->deblock > 2) return fail(OHCU_EINVAL, "bad deblock mode %d", n->deblock);
    j.n_pu = (int)n->n_pu; j.n_tu = (int)n->n_tu; j.n_slices = n->n_slices; j.n_iu = (int)n->n_iu;
    j.ncoeff = (size_t)n16; j.n_coef_units = n->n_coef; j.coef_form = n->n_tu ? n->coef_form : 0;
    j.has_sl = n->scaling_list != 0;
    j.has_deblock = n->deblock != 0; j.bs_derived = n->deblock == OHCU_DEBLOCK_DERIVED;
    j.has_sao = n->sao != 0; j.has_pcm = j.pp.pcm_or_bypass && n->pcm;
    j.all_inter = n->all_inter != 0; j.all_covered = j.all_inter && n->n_pu > 0;
    j.inplace = true;
    return 0;
}

// ---- PCM samples (hls_pcm_sample -> put_pcm): arenas allocated when a stream first uses PCM
extern "C" int ohcu_append_pcm(ohcu_ctx *c, int job, int x0, int y0, int log2_size, int c_idx, const void *samples)
{
    OPEN_JOB(c, job);
    const int n = 1 << log2_size;
    if (!samples || log2_size < 2 || log2_size > 5 || c_idx < 0 || c_idx >= c->nplanes || x0 < 0 || y0 < 0 ||
        x0 + n > c->pw[c_idx] || y0 + n > c->ph[c_idx])
        return fail(OHCU_EINVAL, "PCM block %dx%d at (%d, %d) of plane %d is malformed", n, n, x0, y0, c_idx);
    if (!j.pcm_h) {
        const size_t nsamp = (size_t)c->pw[0] * c->ph[0] + (c->nplanes == 3 ? 2 * (size_t)c->pw[1] * c->ph[1] : 0);
        j.pcm_cap = (int)(nsamp / 16 + 1);
        j.pcms_cap = nsamp * c->P;
        CU(cudaHostAlloc((void **)&j.pcm_h, sizeof(PcmRec) * j.pcm_cap, cudaHostAllocDefault));
        CU(cudaMalloc((void **)&j.pcm_d, sizeof(PcmRec) * j.pcm_cap));
        CU(cudaHostAlloc((void **)&j.pcms_h, j.pcms_cap, cudaHostAllocDefault));
        CU(cudaMalloc((void **)&j.pcms_d, j.pcms_cap));
    }
    const size_t bytes = (size_t)n * n * c->P;
    if (j.n_pcm >= j.pcm_cap || j.pcm_bytes + bytes > j.pcms_cap) return fail(OHCU_EOVERFLOW, "PCM arena full");
    PcmRec r{};
    r.x0 = (uint16_t)x0; r.y0 = (uint16_t)y0; r.log2_size = (uint8_t)log2_size; r.c_idx = (uint8_t)c_idx; r.off = (uint32_t)j.pcm_bytes;
    j.pcm_h[j.n_pcm++] = r;
    memcpy(j.pcms_h + j.pcm_bytes, samples, bytes);
    j.pcm_bytes += (bytes + 15) & ~(size_t)15;
    return 0;
}

// ------------------------------------------------------------------ stage / launch
// PUs of one CTB must be contiguous for the CTB-per-CTA MC kernel.  Decoders emit them
// that way (hls_decode_entry walks CTBs); other orders get a stable counting sort here.
static void sort_pus_by_ctb(ohcu_ctx *c, Job &j)
{
    const int l2 = c->sp.log2_ctb_size;
    std::vector<ohcu_pu> tmp(j.pus.h, j.pus.h + j.n_pu);
    int2 *range = j.ctb_range.h;
    int acc = 0;
    for (int i = 0; i < c->nctb; i++) { range[i].x = acc; acc += range[i].y; range[i].y = 0; }
    for (const ohcu_pu &pu : tmp) {
        const int ctb = (pu.x0 >> l2) + (pu.y0 >> l2) * c->ctb_w;
        j.pus.h[range[ctb].x + range[ctb].y++] = pu;
    }
    j.pu_sorted = true;
}

// Same for the TBs of one CTB (fused MC + residual runs one CTA per CTB).  cross_tu
// references are frame-list indices and follow their block.
static void sort_tus_by_ctb(ohcu_ctx *c, Job &j)
{
    const int l2 = c->sp.log2_ctb_size;
    std::vector<ohcu_tu> tmp(j.tus.h, j.tus.h + j.n_tu);
    std::vector<uint32_t> new_index(j.n_tu);
    int2 *range = j.ctb_tu_range.h;
    int acc = 0;
    for (int i = 0; i < c->nctb; i++) { range[i].x = acc; acc += range[i].y; range[i].y = 0; }
    for (int i = 0; i < j.n_tu; i++) {
        const ohcu_tu &tu = tmp[i];
        const int ths = tu.c_idx ? c->hs : 0, tvs = tu.c_idx ? c->vs : 0;
        const int ctb = ((tu.x0 << ths) >> l2) + ((tu.y0 << tvs) >> l2) * c->ctb_w;
        new_index[i] = (uint32_t)(range[ctb].x + range[ctb].y++);
        j.tus.h[new_index[i]] = tu;
    }
    for (int i = 0; i < j.n_tu; i++)
        if ((j.tus.h[i].flags & OHCU_TUF_CROSS) && j.tus.h[i].cross_tu < (uint32_t)j.n_tu)
            j.tus.h[i].cross_tu = new_index[j.tus.h[i].cross_tu];
    j.tu_sorted = true;
}

static int stage_impl(ohcu_ctx *c, int job, bool close);
static void fill_frame_dev(const ohcu_ctx *c, const Job &j, FrameDev *f);
// A stage that fails discards the job's lists (the job can be begun again).
extern "C" int ohcu_frame_stage(ohcu_ctx *c, int job)
{
    const int rc = stage_impl(c, job, true);
    if (rc && c && job >= 0 && job < (int)c->jobs.size()) c->jobs[job].open = false;
    return rc;
}
// close = false: the job stays the caller's (ohcu_frame_begin keeps refusing it) until ohcu_frame_flush has enqueued
// its launches; closing at the end of the upload let another frame thread begin the job - and reset its lists -
// between the two halves of a flush once the ring wrapped around.
static int stage_impl(ohcu_ctx *c, int job, bool close)
{
    OPEN_JOB(c, job);
    if (!j.inplace && !j.pu_sorted) sort_pus_by_ctb(c, j);
    if (!j.inplace && !j.tu_sorted) sort_tus_by_ctb(c, j);
    // CTBs whose PUs cover all of their samples need no read of the picture before MC
    if (!j.inplace) {
        const int cs = 1 << c->sp.log2_ctb_size;
        j.all_covered = j.n_pu > 0;
        for (int y = 0; y < c->ctb_h; y++)
            for (int x = 0; x < c->ctb_w; x++) {
                const int rs = y * c->ctb_w + x;
                const int area = std::min(cs, c->sp.width - x * cs) * std::min(cs, c->sp.height - y * cs);
                int2 &r = j.ctb_range.h[rs];
                r.y &= ~CTB_COVERED;
                if (j.ctb_area[rs] > area || (r.y & 0xffffff) > 128)
                    return fail(OHCU_EINVAL, "CTB %d: PUs overlap or exceed the CTB (%d PUs, %d samples)", rs, r.y & 0xffffff,
                                j.ctb_area[rs]);
                if (j.ctb_area[rs] == area) r.y |= CTB_COVERED;
                else j.all_covered = false;
            }
    }
    if (!j.inplace && j.max_slice >= j.n_slices) return fail(OHCU_EINVAL, "a PU uses slice %d which was never set", j.max_slice);
    // every reference a PU names must be a DPB slot
    for (int i = 0; !j.inplace && i < j.n_pu; i++) {
        const ohcu_pu &pu = j.pus.h[i];
        for (int l = 0; l < 2; l++)
            if (pu.pred_flag & (1 << l)) {
                const int slot = j.slices.h[pu.slice].ref_slot[l][pu.ref_idx[l]];
                if (slot >= (int)c->slots.size())
                    return fail(OHCU_EINVAL, "PU record %d: list %d, ref_idx %d names DPB slot %d (of %zu)", i, l, pu.ref_idx[l], slot, c->slots.size());
            }
    }
    for (int i = 0; !j.inplace && j.has_cross && i < j.n_tu; i++)
        if ((j.tus.h[i].flags & OHCU_TUF_CROSS) &&
            (j.tus.h[i].cross_tu >= (uint32_t)j.n_tu || j.tus.h[j.tus.h[i].cross_tu].log2_size != j.tus.h[i].log2_size))
            return fail(OHCU_EINVAL, "TU %d: bad cross-component reference", i);
    cudaStream_t st = j.stream;
    CU(push_arena(j.pus, j.n_pu, st));
    CU(push_arena(j.tus, j.n_tu, st));
    CU(push_arena(j.coeffs, j.ncoeff, st));
    if (j.coef_form == 3) {
        for (int i = 0; !j.inplace && !j.has_sl && i < j.n_tu; i++)
            if (j.tuq.h[i].matrix_id != OHCU_QM_FLAT) return fail(OHCU_EINVAL, "TU %d names a scaling list but none was set", i);
        CU(push_arena(j.tuq, j.n_tu, st));
        if (j.has_sl) CU(cudaMemcpyAsync(j.sl_d, j.sl_h, SL_BYTES, cudaMemcpyHostToDevice, st));
    }
    // the tables: one copy up to the last group this frame uses
    const size_t tab_bytes = j.tab_end[j.has_pcm ? 4 : j.has_deblock ? (j.bs_derived ? 2 : 3) : j.has_sao ? 1 : 0];
    CU(cudaMemcpyAsync(j.tab_d, j.tab_h, tab_bytes, cudaMemcpyHostToDevice, st));
    if (j.n_pcm) {
        CU(cudaMemcpyAsync(j.pcm_d, j.pcm_h, sizeof(PcmRec) * j.n_pcm, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(j.pcms_d, j.pcms_h, j.pcm_bytes, cudaMemcpyHostToDevice, st));
    }
    if (j.n_iu) {
        CU(push_arena(j.ius, j.n_iu, st));
        CU(push_arena(j.iu_ctbs, c->nctb, st));
    }
    CU(cudaMemsetAsync(j.err_d, 0, sizeof(int), st));
    if (j.inplace) {
        // validation, per-CTB ranges and flags on the device (prepare.cu)
        FrameDev f;
        fill_frame_dev(c, j, &f);
        PrepDev p{};
        p.idx = j.ctbs.d; p.pu_range = j.ctb_range.d; p.tu_range = j.ctb_tu_range.d; p.coef_range = j.ctb_coef_range.d;
        p.sao_flags = j.sao_flags.d; p.bs_flags = j.bs_flags.d;
        p.fse = j.fse.d; p.tsa = j.tsa.d; p.tile = j.tile.d; p.dbk_disabled = j.dbk_disabled.d;
        p.err = j.err_d;
        p.n_pu = (uint32_t)j.n_pu; p.n_tu = (uint32_t)j.n_tu; p.n_coef = j.n_coef_units;
        p.n_slices = j.n_slices; p.coef_form = j.coef_form; p.dpb_slots = (int)c->slots.size(); p.out_slot = j.out_slot;
        p.all_inter = j.all_inter; p.no_tile_filter = j.pp.no_tile_filter; p.has_sao = j.has_sao; p.derive = j.bs_derived;
        p.area = recon_sparse_ctb_coefs(c->sp.log2_ctb_size, c->nplanes, c->hs, c->vs); p.has_sl = j.has_sl;
        launch_prepare(f, p, st);
        CU(cudaGetLastError());
        std::lock_guard<std::mutex> lock(c->mu);
        c->launches++;
    }
    if (close) {
        std::lock_guard<std::mutex> lock(c->mu);
        j.open = false;
    }
    j.staged = true;
    // the pinned staging is reusable once this event has fired
    CU(cudaEventRecord(j.done, st));
    j.has_done = true;
    return 0;
}

static void fill_frame_dev(const ohcu_ctx *c, const Job &j, FrameDev *f)
{
    fill_seq_dev(c, f);
    f->pus = j.pus.d; f->n_pu = j.n_pu; f->ctb_pu_range = j.ctb_range.d; f->slices = j.slices.d;
    f->tus = j.tus.d; f->n_tu = j.n_tu; f->coeffs = j.coeffs.d; f->ctb_tu_range = j.ctb_tu_range.d;
    f->ctb_coef_range = j.coef_form >= 2 ? j.ctb_coef_range.d : nullptr;
    f->tu_quant = j.coef_form == 3 ? j.tuq.d : nullptr;
    f->scaling = j.coef_form == 3 && j.has_sl ? j.sl_d : nullptr;
    f->vbs = j.has_deblock ? j.vbs.d : nullptr; f->hbs = j.hbs.d; f->qp_y = j.qp.d; f->db = j.db.d;
    f->is_pcm = j.has_pcm ? j.is_pcm.d : nullptr;
    f->bs_ctb_flags = j.has_deblock && j.bs_derived ? j.bs_flags.d : nullptr;
    f->sao = j.has_sao ? j.sao.d : nullptr; f->sao_ctb_flags = j.sao_flags.d;
    f->pcm_or_bypass = j.pp.pcm_or_bypass; f->cb_qp_offset = j.pp.cb_qp_offset; f->cr_qp_offset = j.pp.cr_qp_offset;
    f->ius = j.ius.d; f->iu_index = j.iu_ctbs.d; f->n_iu = j.n_iu;
    for (int i = 0; i < 3; i++) { f->resid[i] = (int16_t *)j.resid.p[i]; f->resid_pitch[i] = j.resid.pitch[i]; }
    f->intra_progress = j.progress_d; f->intra_order = c->intra_order_d;
    f->intra_smoothing_disabled = c->sp.intra_smoothing_disabled; f->strong_intra_smoothing = c->sp.strong_intra_smoothing;
    f->err = j.err_d;
}

// stages: bit mask 1 MC, 2 residual, 4 deblock, 8 SAO (0 = all).  Test / bench hook behind
// ohcu_frame_launch; when MC is excluded the slot's current content is the starting picture.
extern "C" int ohcu_frame_launch_stages(ohcu_ctx *c, int job, int out_slot, int stages)
{
    ENTER(c);
    if (check_job(c, job)) return OHCU_EINVAL;
    Job &j = c->jobs[job];
    if (!j.staged) return fail(OHCU_ESTATE, "job %d has no staged lists", job);
    if (out_slot < 0) out_slot = j.out_slot;
    if (check_slot(c, out_slot)) return OHCU_EINVAL;
    if (!stages) stages = 15;
    cudaStream_t st = j.stream;
    Slot &os = c->slots[out_slot];
    std::lock_guard<std::mutex> lock(c->mu);

    // ---- ordering between frames: replaces ff_thread_await_progress (hevcdec.c:2051-2058).
    // Readers: every slot the slice tables name.  Writer: wait for earlier readers/downloads.
    uint32_t ref_mask = 0;
    if (stages & 1)
        for (int s = 0; s < j.n_slices; s++)
            for (int l = 0; l < 2; l++)
                for (int r = 0; r < OHCU_MAX_REFS; r++) {
                    const int slot = j.slices.h[s].ref_slot[l][r];
                    if (slot < (int)c->slots.size()) ref_mask |= 1u << slot;
                }
    if (ref_mask & (1u << out_slot)) ref_mask &= ~(1u << out_slot);   // a picture never references itself
    for (size_t s = 0; s < c->slots.size(); s++)
        if ((ref_mask >> s) & 1) {
            if (c->slots[s].has_ready) CU(cudaStreamWaitEvent(st, c->slots[s].ready, 0));
            c->slots[s].readers |= 1ull << job;
        }
    for (size_t k = 0; k < c->jobs.size(); k++)
        if (((os.readers >> k) & 1) && (int)k != job && c->jobs[k].has_done)
            CU(cudaStreamWaitEvent(st, c->jobs[k].done, 0));
    os.readers = 0;
    if (os.has_ready) CU(cudaStreamWaitEvent(st, os.ready, 0));
    if (os.has_read) CU(cudaStreamWaitEvent(st, os.read_done, 0));

    FrameDev f;
    fill_frame_dev(c, j, &f);
    const bool do_sao = c->sp.sao_enabled && j.has_sao && (stages & 8);
    const int lf_flags = (((stages & 4) && j.has_deblock) ? 1 : 0) | (do_sao ? 2 : 0);
    f.out = os.pic.ps;
    // with a loop filter to follow, prediction + residual go to the job's own picture and the
    // filter kernel writes the slot; otherwise they write the slot directly
    f.work = lf_flags ? j.recon.ps : os.pic.ps;
    const bool timed = c->profiling;
    int n, total = 0;
#define STAGE_EVENT(i) do { if (timed) CU(cudaEventRecord(j.ev[i], st)); } while (0)
#define LAUNCHED(call) do { n = (call); if (n < 0) return fail(OHCU_EINVAL, "launch refused"); total += n; \
        CU(cudaGetLastError()); } while (0)
    STAGE_EVENT(0);
    // boundary strengths derived from the lists (bs.cu): on the job's side stream, beside mc_kernel and residual_kernel;
    // it starts after everything already queued on the job's stream (staging copies, the previous launch)
    const bool derive = (lf_flags & 1) && f.bs_ctb_flags;
    if (derive) {
        CU(cudaEventRecord(j.fork, st));
        CU(cudaStreamWaitEvent(j.side, j.fork, 0));
        LAUNCHED(launch_bs_derive(f, j.side));
        CU(cudaEventRecord(j.join, j.side));
    }
    // samples this launch does not predict keep the slot's current content
    if ((!(stages & 1) || !j.all_covered) && lf_flags) LAUNCHED(launch_copy_picture(f, f.work, f.out, st));
    // prediction (mc_kernel), then the coded blocks on top of it (residual_kernel), recon.cu
    LAUNCHED(launch_recon(f, stages & 1, st));
    STAGE_EVENT(1);
    LAUNCHED(launch_recon(f, stages & 2, st));
    // PCM blocks replace prediction + residual of their samples (put_pcm), ahead of the loop filters
    if ((stages & 2) && j.n_pcm) LAUNCHED(launch_pcm(f, j.pcm_d, j.n_pcm, j.pcms_d, st));
    // the intra blocks, predicted from what has been reconstructed around them, with their residual (intra.cu)
    if ((stages & 2) && j.n_iu) {
        CU(cudaMemsetAsync(j.progress_d, 0, sizeof(int) * (c->nctb + 1), st));
        LAUNCHED(launch_intra(f, st));
    }
    STAGE_EVENT(2);
    // deblocking and SAO run fused (loopfilter.cu); stage 3 carries their time
    if (derive) CU(cudaStreamWaitEvent(st, j.join, 0));
    if (lf_flags) LAUNCHED(launch_loopfilter(f, j.lf_maps, lf_flags, st));
    STAGE_EVENT(3);
    // the finished picture becomes a reference: replicate its borders into the halo
    LAUNCHED(launch_pad_picture(f, f.out, st));
    STAGE_EVENT(4);
#undef STAGE_EVENT
#undef LAUNCHED
    j.timed = timed;
    c->launches += total;
    // the job's error word travels back with its results (read by ohcu_frame_sync / ohcu_slot_sync)
    if (j.inplace) CU(cudaMemcpyAsync(j.err_h, j.err_d, sizeof(int), cudaMemcpyDeviceToHost, st));
    os.producer = job; os.released = false;
    CU(cudaEventRecord(j.done, st));
    j.has_done = true;
    CU(cudaEventRecord(os.ready, st));
    os.has_ready = true;
    return 0;
}

extern "C" int ohcu_frame_launch(ohcu_ctx *c, int job, int out_slot)
{
    return ohcu_frame_launch_stages(c, job, out_slot, 0);
}

extern "C" int ohcu_frame_flush(ohcu_ctx *c, int job)
{
    int rc = stage_impl(c, job, false);
    if (!rc) rc = ohcu_frame_launch_stages(c, job, -1, 0);
    if (c && job >= 0 && job < (int)c->jobs.size()) {
        std::lock_guard<std::mutex> lock(c->mu);
        c->jobs[job].open = false;
    }
    return rc;
}

extern "C" int ohcu_frame_sync(ohcu_ctx *c, int job)
{
    ENTER(c);
    if (check_job(c, job)) return OHCU_EINVAL;
    CU(cudaStreamSynchronize(c->jobs[job].stream));
    return job_list_error(c, job);
}

extern "C" int ohcu_sync_all(ohcu_ctx *c)
{
    ENTER(c);
    CU(cudaDeviceSynchronize());
    return 0;
}

// ------------------------------------------------------------------ measurement hooks
extern "C" int ohcu_set_profiling(ohcu_ctx *c, int enable)
{
    if (!c) return fail(OHCU_EINVAL, "null context");
    c->profiling = enable != 0;
    return 0;
}

extern "C" int ohcu_job_stage_ms(ohcu_ctx *c, int job, float ms[OHCU_NUM_STAGES])
{
    ENTER(c);
    if (check_job(c, job)) return OHCU_EINVAL;
    Job &j = c->jobs[job];
    if (!j.timed) return fail(OHCU_ESTATE, "job %d was not launched with profiling enabled", job);
    CU(cudaEventSynchronize(j.ev[4]));
    for (int i = 0; i < 4; i++) CU(cudaEventElapsedTime(&ms[i], j.ev[i], j.ev[i + 1]));
    CU(cudaEventElapsedTime(&ms[OHCU_STAGE_TOTAL], j.ev[0], j.ev[4]));
    return 0;
}

extern "C" int ohcu_job_read_bs(ohcu_ctx *c, int job, uint8_t *vertical_bs, uint8_t *horizontal_bs)
{
    ENTER(c);
    if (check_job(c, job)) return OHCU_EINVAL;
    Job &j = c->jobs[job];
    if (!vertical_bs || !horizontal_bs) return fail(OHCU_EINVAL, "null argument");
    if (!j.staged || !j.has_deblock) return fail(OHCU_ESTATE, "job %d holds no deblocking input", job);
    CU(cudaStreamSynchronize(j.stream));
    CU(cudaMemcpy(vertical_bs, j.vbs.d, (size_t)c->bs_w * c->bs_h, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(horizontal_bs, j.hbs.d, (size_t)c->bs_w * c->bs_h, cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int64_t ohcu_kernel_launches(const ohcu_ctx *c) { return c ? c->launches : 0; }

extern "C" void *ohcu_job_stream(ohcu_ctx *c, int job)
{
    if (!c || job < 0 || job >= (int)c->jobs.size()) return nullptr;
    return (void *)c->jobs[job].stream;
}

extern "C" int ohcu_flush_l2(ohcu_ctx *c)
{
    ENTER(c);
    if (!c->l2_buf) {
        c->l2_bytes = (size_t)256 << 20;          // > 126 MB L2
        CU(cudaMalloc(&c->l2_buf, c->l2_bytes));
        CU(cudaMemset(c->l2_buf, 0, c->l2_bytes));
    }
    CU(cudaDeviceSynchronize());
    launch_l2_flush(c->l2_buf, c->l2_bytes, 0);
    CU(cudaGetLastError());
    CU(cudaDeviceSynchronize());
    return 0;
}

extern "C" int ohcu_timer_start(ohcu_ctx *c)
{
    ENTER(c);
    if (!c->t0) { CU(cudaEventCreate(&c->t0)); CU(cudaEventCreate(&c->t1)); }
    CU(cudaDeviceSynchronize());
    CU(cudaEventRecord(c->t0, c->copy_stream));
    // nothing enqueued afterwards may start before the start event
    for (auto &j : c->jobs) CU(cudaStreamWaitEvent(j.stream, c->t0, 0));
    return 0;
}

extern "C" int ohcu_timer_stop(ohcu_ctx *c, float *ms)
{
    ENTER(c);
    if (!c->t0 || !ms) return fail(OHCU_ESTATE, "timer not started");
    for (auto &j : c->jobs)
        if (j.has_done) CU(cudaStreamWaitEvent(c->copy_stream, j.done, 0));
    for (auto &s : c->slots)
        if (s.has_read) CU(cudaStreamWaitEvent(c->copy_stream, s.read_done, 0));
    CU(cudaEventRecord(c->t1, c->copy_stream));
    CU(cudaEventSynchronize(c->t1));
    CU(cudaEventElapsedTime(ms, c->t0, c->t1));
    return 0;
}

extern "C" int ohcu_host_alloc(size_t bytes, void **ptr)
{
    if (!ptr) return fail(OHCU_EINVAL, "null argument");
    CU(cudaHostAlloc(ptr, bytes ? bytes : 1, cudaHostAllocDefault));
    return 0;
}

extern "C" int ohcu_host_free(void *ptr)
{
    if (ptr) CU(cudaFreeHost(ptr));
    return 0;
}

// ------------------------------------------------------------------ per-entry test shims
// One block through the device code of the frame kernels, with the argument meaning of the
// HEVCDSPContext entries (hevcdsp.h:51-166).  Host pointers; synchronous; tests only.
#define DSP_ENTER()                                                                      \
    do {                                                                                 \
        int n_ = 0;                                                                      \
        if (cudaGetDeviceCount(&n_) != cudaSuccess || n_ < 1) return fail(OHCU_ENODEV, "no CUDA device"); \
    } while (0)
#define DSP_RC(call) ((call) == 0 ? 0 : fail(OHCU_ECUDA, "per-entry shim failed: %s", cudaGetErrorString(cudaGetLastError())))

extern "C" int ohcu_dsp_mc_put(int bd, int luma, int16_t *dst, const uint8_t *src, ptrdiff_t srcstride, int height, int mx, int my, int width)
{
    DSP_ENTER();
    return DSP_RC(dsp_mc(bd, luma, 0, nullptr, 0, src, srcstride, nullptr, dst, height, 0, 0, 0, 0, 0, mx, my, width));
}
extern "C" int ohcu_dsp_mc_uni(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src, ptrdiff_t srcstride, int height,
                               int mx, int my, int width)
{
    DSP_ENTER();
    return DSP_RC(dsp_mc(bd, luma, 1, dst, dststride, src, srcstride, nullptr, nullptr, height, 0, 0, 0, 0, 0, mx, my, width));
}
extern "C" int ohcu_dsp_mc_uni_w(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src, ptrdiff_t srcstride, int height,
                                 int denom, int wx, int ox, int mx, int my, int width)
{
    DSP_ENTER();
    return DSP_RC(dsp_mc(bd, luma, 2, dst, dststride, src, srcstride, nullptr, nullptr, height, denom, wx, 0, ox, 0, mx, my, width));
}
extern "C" int ohcu_dsp_mc_bi(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src, ptrdiff_t srcstride,
                              const int16_t *src2, int height, int mx, int my, int width)
{
    DSP_ENTER();
    return DSP_RC(dsp_mc(bd, luma, 3, dst, dststride, src, srcstride, src2, nullptr, height, 0, 0, 0, 0, 0, mx, my, width));
}
extern "C" int ohcu_dsp_mc_bi_w(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src, ptrdiff_t srcstride,
                                const int16_t *src2, int height, int denom, int wx0, int wx1, int ox0, int ox1, int mx, int my, int width)
{
    DSP_ENTER();
    return DSP_RC(dsp_mc(bd, luma, 4, dst, dststride, src, srcstride, src2, nullptr, height, denom, wx0, wx1, ox0, ox1, mx, my, width));
}
extern "C" int ohcu_dsp_idct(int bd, int16_t *coeffs, int log2_size, int col_limit)
{
    DSP_ENTER();
    return DSP_RC(dsp_residual(bd, OHCU_TU_IDCT, 0, coeffs, log2_size, col_limit, 1));
}
extern "C" int ohcu_dsp_idct_dc(int bd, int16_t *coeffs, int log2_size)
{
    DSP_ENTER();
    return DSP_RC(dsp_residual(bd, OHCU_TU_DC, 0, coeffs, log2_size, 0, 0));
}
extern "C" int ohcu_dsp_idct_4x4_luma(int bd, int16_t *coeffs)
{
    DSP_ENTER();
    return DSP_RC(dsp_residual(bd, OHCU_TU_DST4, 0, coeffs, 2, 0, 0));
}
extern "C" int ohcu_dsp_transform_skip(int bd, int16_t *coeffs, int log2_size)
{
    DSP_ENTER();
    return DSP_RC(dsp_residual(bd, OHCU_TU_SKIP, 0, coeffs, log2_size, 0, 0));
}
extern "C" int ohcu_dsp_transform_rdpcm(int bd, int16_t *coeffs, int log2_size, int mode)
{
    DSP_ENTER();
    return DSP_RC(dsp_residual(bd, OHCU_TU_BYPASS, OHCU_TUF_RDPCM | (mode ? OHCU_TUF_RDPCM_VERT : 0), coeffs, log2_size, 0, 0));
}
extern "C" int ohcu_dsp_transform_add(int bd, uint8_t *dst, const int16_t *coeffs, ptrdiff_t stride, int n)
{
    DSP_ENTER();
    return DSP_RC(dsp_transform_add(bd, dst, coeffs, stride, n));
}
extern "C" int ohcu_dsp_intra_pred(int bd, uint8_t *plane, ptrdiff_t stride, int pw, int ph, int x, int y, int log2_size, int c_idx, int mode,
                                   int cand, int filter_ok, int strong)
{
    DSP_ENTER();
    if (!plane || (bd != 8 && bd != 9 && bd != 10 && bd != 12 && bd != 14)) return fail(OHCU_EINVAL, "bad plane or bit depth");
    return DSP_RC(dsp_intra_pred(bd, plane, stride, pw, ph, x, y, log2_size, c_idx, mode, cand, filter_ok, strong));
}
extern "C" int ohcu_dsp_loop_filter_luma(int bd, int vertical, uint8_t *pix, ptrdiff_t stride, int beta, const int *tc, const uint8_t *no_p,
                                         const uint8_t *no_q)
{
    DSP_ENTER();
    return DSP_RC(dsp_loop_filter(bd, 1, vertical, pix, stride, beta, tc, no_p, no_q));
}
extern "C" int ohcu_dsp_loop_filter_chroma(int bd, int vertical, uint8_t *pix, ptrdiff_t stride, const int *tc, const uint8_t *no_p,
                                           const uint8_t *no_q)
{
    DSP_ENTER();
    return DSP_RC(dsp_loop_filter(bd, 0, vertical, pix, stride, 0, tc, no_p, no_q));
}
extern "C" int ohcu_dsp_sao_band_filter(int bd, uint8_t *dst, const uint8_t *src, ptrdiff_t stride_dst, ptrdiff_t stride_src,
                                        const int16_t *offset_val, int band_position, int width, int height)
{
    DSP_ENTER();
    return DSP_RC(dsp_sao(bd, 0, dst, src, stride_dst, stride_src, offset_val, band_position, width, height));
}
extern "C" int ohcu_dsp_sao_edge_filter(int bd, uint8_t *dst, const uint8_t *src, ptrdiff_t stride_dst, ptrdiff_t stride_src,
                                        const int16_t *offset_val, int eo, int width, int height)
{
    DSP_ENTER();
    return DSP_RC(dsp_sao(bd, 1, dst, src, stride_dst, stride_src, offset_val, eo, width, height));
}
