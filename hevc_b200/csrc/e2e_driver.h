/*
 * e2e_driver.h — interface of the plain-C multi-threaded caller of include/ohcu.h (e2e_driver.c).
 * Measurement / test harness, not part of the product ABI: bench.py's e2e leg and tests/test_gpu_driver.py
 * bind it through ctypes (hevc_b200/driver.py).  Built as hevc_b200/libohcu_drv.so, linked against libohcu.so.
 */
#ifndef OHCU_E2E_DRIVER_H
#define OHCU_E2E_DRIVER_H

#include <stddef.h>
#include <stdint.h>

#include "../../include/ohcu.h"

#ifdef __cplusplus
extern "C" {
#endif

#define OHDRV_MAX_THREADS 64

/* how a frame thread hands a frame's lists to the library */
#define OHDRV_FILL_COPY 0   /* in-place path; record lists + coefficients memcpy'd into the job's arenas inside the
                               timed region (lower bound of the stores a parser makes while it decodes)        */
#define OHDRV_FILL_NONE 1   /* in-place path; the big lists (PUs, TUs, coefficients) are found in the arena where an
                               earlier pass left them, only the small tables are written (the library alone)   */
#define OHDRV_APPEND    2   /* the copying, host-validated ohcu_append_* / ohcu_set_* entry points             */

#define OHDRV_DOWNLOAD_NONE    0
#define OHDRV_DOWNLOAD_PLANES  1   /* ohcu_slot_download_async into out_planes[3*k..] with out_strides            */
#define OHDRV_DOWNLOAD_CROPPED 2   /* ohcu_slot_output_cropped_async (full window) into out_planes[3*k..]          */

/* one frame's work lists, prepared ahead (borrowed pointers) */
typedef struct ohdrv_frame {
    const ohcu_pu *pus;             uint32_t n_pu;
    uint32_t n_tu;
    const ohcu_tu *tus;             /* sparse / levels: records as ohcu_append_tu_sparse takes them            */
    const ohcu_tu_quant *quant;     /* levels form, else NULL                                                  */
    const void *coefs;              /* ohcu_coef pairs (sparse / levels) or int16 (dense)                      */
    uint64_t n_coef;                /* pairs or int16                                                          */
    int32_t coef_form;              /* OHCU_COEF_*                                                             */
    int32_t n_slices;
    const ohcu_slice_tab *slices;
    const ohcu_ctb_index *ctbs;     /* per CTB (in-place path)                                                 */
    int32_t deblock;                /* OHCU_DEBLOCK_*                                                          */
    int32_t has_sao;
    const uint8_t *vbs, *hbs;
    const int8_t *qp;
    const ohcu_dbparams *db;
    const uint8_t *is_pcm;          /* or NULL                                                                 */
    const ohcu_sao *sao;
    const uint8_t *fse;
    const int32_t *tsa, *tile;
    const uint8_t *scaling_list;    /* 1548 bytes or NULL                                                      */
    int32_t all_inter;
    uint32_t n_iu;                  /* intra blocks (the context needs ohcu_seq_params.intra)                  */
    const ohcu_iu *ius;
    const ohcu_iu_index *iu_ctbs;   /* per CTB                                                                 */
} ohdrv_frame;

typedef struct ohdrv_config {
    int32_t threads;                /* frame threads (hevcdec.h:53: the reference runs up to 16)               */
    int32_t total_frames;
    int32_t ring;                   /* output slots / pinned output pictures in rotation (<= 64)               */
    int32_t first_slot;             /* ring entry k writes DPB slot first_slot + k                             */
    int32_t fill;                   /* OHDRV_FILL_* / OHDRV_APPEND                                             */
    int32_t download;               /* OHDRV_DOWNLOAD_*                                                        */
    int32_t timed;                  /* bracket the run with ohcu_timer_start / ohcu_timer_stop                 */
    int32_t chain_streams;          /* 0: frames are independent (their slice tables name the references).
                                       S > 0: S independent streams; picture t of a stream predicts from the
                                       stream's pictures t-1 and t-2 (slots rotate over ring / S >= 3 entries) */
    int32_t ref_slot[2];            /* chain mode: what the first two pictures of a stream predict from        */
    int32_t frame_by_job;           /* the frame a job carries is frames[job id % n_frames] instead of the frame
                                       counter's: with OHDRV_FILL_NONE the big lists an earlier OHDRV_FILL_COPY
                                       pass left in the job's arena are then the right ones                       */
    int32_t pad;
    int32_t *last_frame;            /* optional, ring entries: index of the frame whose picture the entry holds   */
    uint8_t *const *out_planes;     /* ring * 3 plane pointers (pinned), or NULL                               */
    const ptrdiff_t *out_strides;   /* 3 strides (OHDRV_DOWNLOAD_PLANES)                                       */
} ohdrv_config;

typedef struct ohdrv_result {
    float device_ms;                /* ohcu_timer_stop: device time from the start event to the last job / download */
    double wall_ms;                 /* host wall clock of the same region                                      */
    double host_ms;                 /* until the last frame thread returned (everything enqueued)              */
    double thread_cpu_ms;           /* CPU time summed over the frame threads (includes CUDA's spin-waits)     */
    double work_cpu_ms;             /* of which: filling the lists, flush, download calls (no waiting)         */
    int32_t frames;
    char error[512];
} ohdrv_result;

int ohdrv_run(ohcu_ctx *ctx, const ohdrv_frame *frames, int n_frames, const ohdrv_config *cfg, ohdrv_result *res);

#ifdef __cplusplus
}
#endif
#endif
