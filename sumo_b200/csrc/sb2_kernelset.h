/* sb2_kernelset.h -- table of launchers one kernel set (kernels.cu compiled with one feature selection) exports to the
 * engine.  `ctx` points at the engine's sb2::Ctx; the launchers pass it to their kernels by value. */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/sumob200.h"

struct sb2_kernelset {
    const char* name;
    int block;                 // threads per CTA of every kernel
    size_t ctx_size;           // sizeof(Ctx) the set was compiled with (checked by the engine)
    void (*reset)(const void* ctx, cudaStream_t s);
    void (*pre)(const void* ctx, int n_threads, cudaStream_t s);
    void (*plan)(const void* ctx, int n_vehicles, cudaStream_t s);
    void (*execute)(const void* ctx, int n_vehicles, cudaStream_t s);
    void (*remove)(const void* ctx, int n_lanes, cudaStream_t s);
    void (*integrate)(const void* ctx, int n_lanes, cudaStream_t s);
    void (*insert)(const void* ctx, int n_lanes, cudaStream_t s);
    void (*count)(const void* ctx, int n_vehicles, unsigned long long* updates, cudaStream_t s);
    void (*mirror)(const void* ctx, int n_vehicles, sumob200_vehicle_state* out, cudaStream_t s);
    void (*fcd)(const void* ctx, int n_vehicles, int32_t* block_counts /* [blocks + 1] */, sumob200_fcd_row* rows, cudaStream_t s);   // 3 launches
    void (*tripinfo)(const void* ctx, int n_vehicles, sumob200_tripinfo* out, cudaStream_t s);
    /* summary aggregates (sb2_summary.h:SummaryPart): parts == NULL returns the scratch size in bytes, else launches (returns the launch count) */
    int (*summary)(const void* ctx, int n_vehicles, void* parts, void* out, cudaStream_t s);
    void (*tls_set_phase)(const void* ctx, int tls, int phase, cudaStream_t s);
    void (*set_speed)(const void* ctx, const int32_t* slots, const double* speeds, int n, cudaStream_t s);
    /* halo exchange of the edge partition (sb2_halo.h); `side` points at an sb2::HaloSide in host memory, msg at device memory */
    int (*halo_pack)(const void* ctx, const void* side, const uint8_t* imp_mask, void* msg, cudaStream_t s);      /* returns the launch count */
    int (*halo_unpack)(const void* ctx, const void* side, const void* msg, cudaStream_t s);
    /* re-slotting (engine.cu reslot): the permutation of the on-road vehicles into lane order, and the fix-up of every stored slot number */
    int (*reslot_maps)(const void* ctx, int n_lanes, int n_vehicles, int32_t* blocks_a, int32_t* blocks_b, int32_t* order, int32_t* sorted,
                       int32_t* new_of_old, int32_t* old_of_new, cudaStream_t s);
    int (*reslot_remap)(const void* ctx, int n_lanes, int n_vehicles, const int32_t* new_of_old, int app_buf, cudaStream_t s);
    /* E1 detectors (sb2_e1.h): append the vehicles that are on a detector right now to the event list */
    void (*e1_collect)(const void* ctx, int n_vehicles, cudaStream_t s);
    /* libsumo::Lane::getLastStep* of every lane */
    void (*lane_stats)(const void* ctx, int n_lanes, sumob200_lane_stats* out, cudaStream_t s);
    /* E2 detectors (sb2_e2.h): detectorUpdate after a step; the interval's attributes (sb2::E2Result per detector) + reset */
    void (*e2_update)(const void* ctx, int n_detectors, cudaStream_t s);
    void (*e2_readout)(const void* ctx, int n_detectors, int buf, void* out, cudaStream_t s);
};

extern "C" const sb2_kernelset* sb2_kernelset_full(void);
extern "C" const sb2_kernelset* sb2_kernelset_krauss(void);
extern "C" const sb2_kernelset* sb2_kernelset_plain(void);
