/*
 * replay.h — frame-level CPU replay of one frame's work lists.
 * TEST INFRASTRUCTURE ONLY (see hevc_oracle.h for the rule).
 *
 * A frame job is exactly what the decoder hands the C ABI (include/ohcu.h) for
 * one picture, plus host pictures for the DPB.  Two replays consume it:
 *   or_replay_frame()    oracle/replay.c   — pure restatement (hevc_oracle.c leaves)
 *   ohref_replay_frame() oracle/ref_shim.c — the reference's OWN code: its C
 *                        HEVCDSPContext table, its luma_mc_* / chroma_mc_* drivers
 *                        (hevcdec.c:1744-2049), and its ff_hevc_hls_filters driver
 *                        (hevc_filter.c:264-762, 1211-1245); only built when
 *                        /root/reference is present (oracle/Makefile -> oracle/_ref/).
 * Both process in the reference's order: all PUs (hls_prediction_unit order),
 * all TUs (transform_add), then per CTB in raster order the one-CTB-delayed
 * deblock + SAO of ff_hevc_hls_filters (hevcdec.c:2898-2903).
 */
#ifndef OH_REPLAY_H
#define OH_REPLAY_H

#include "../include/ohcu.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oh_picture {
    uint8_t  *data[3];
    ptrdiff_t stride[3];       /* bytes */
} oh_picture;

typedef struct oh_frame_job {
    ohcu_seq_params sp;
    int32_t n_slices;
    const ohcu_slice_tab *slices;
    int32_t n_pu;
    const ohcu_pu *pus;
    int32_t n_tu;
    const ohcu_tu *tus;
    const int16_t *coeffs;        /* arena; never modified (blocks are copied) */
    int64_t ncoeff;
    const ohcu_deblock_in *deblock;   /* NULL = no deblocking */
    const ohcu_sao_in *sao;           /* NULL = no SAO        */
    const oh_picture *dpb;            /* sp.dpb_slots reference pictures */
    oh_picture out;                   /* reconstructed picture (written) */
    int32_t stages;                   /* bit mask: 1 MC, 2 residual, 4 deblock, 8 SAO; 0 = all */
} oh_frame_job;

#define OH_STAGE_MC       1
#define OH_STAGE_RESIDUAL 2
#define OH_STAGE_DEBLOCK  4
#define OH_STAGE_SAO      8

/* returns 0, or a negative OHCU_E* code for malformed jobs */
int or_replay_frame(const oh_frame_job *job);
int ohref_replay_frame(const oh_frame_job *job);

#ifdef __cplusplus
}
#endif
#endif
