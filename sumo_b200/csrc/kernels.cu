/* kernels.cu -- the sm_100a kernels of the step pipeline, compiled once per KERNEL SET.
 *
 * A kernel set is this file compiled with a feature selection (-DSB2_KSET=<name> plus -DSB2_NO_* switches); the engine
 * (engine.cu) picks the set that matches the scenario when it is created.  Measured reason (profiles/r01_*): k_plan is bound
 * by divergence, instruction fetch and the 64-register allocation, and car-following code that a scenario never executes
 * (the IDM branches for a Krauss-only fleet) costs 10 % of its time just by being compiled in.
 *   full    every implemented feature
 *   krauss  -DSB2_NO_IDM : all vTypes use carFollowModel Krauss
 *   plain   -DSB2_NO_IDM -DSB2_NO_OBSERVERS : Krauss fleet, no E1 / E2 detectors, no edgeData (the move-reminder hooks cost
 *           k_execute 7 % when compiled in and unused)
 * Every set lives in its own namespace, so the inlined device functions of two sets never meet at link time.
 *
 * One simulation step = 8 launches on one stream (no host synchronisation in between).  Every kernel takes the Ctx by
 * value as a __grid_constant__ parameter, so the ~190 array base pointers are read from the constant bank, not from HBM.
 */
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include <float.h>
#include <cstdio>

#include "../../include/sumob200.h"
#include "sb2_kernelset.h"

#ifndef SB2_KSET
#error "compile with -DSB2_KSET=<name>"
#endif
#define SB2_PASTE2(a, b) a##b
#define SB2_PASTE(a, b) SB2_PASTE2(a, b)
#define SB2_STR2(x) #x
#define SB2_STR(x) SB2_STR2(x)

namespace SB2_PASTE(sb2_kset_, SB2_KSET) {

#include "sb2_types.h"
#include "sb2_cf.h"
#include "sb2_veh.h"
#include "sb2_link.h"
#include "sb2_plan.h"
#include "sb2_plan2.h"
#include "sb2_emis.h"
#include "sb2_exec.h"
#include "sb2_lane.h"
#include "sb2_insert.h"
#include "sb2_transfer.h"
#include "sb2_halo.h"
#include "sb2_fcd.h"
#include "sb2_summary.h"

using namespace sb2;

constexpr int BLOCK = 128;
// resident CTAs per SM the register allocation of the two per-vehicle kernels is bounded for (profiles/r01_*: both are
// latency bound -- long scoreboard on dependent gathers -- so occupancy is worth more than a spill-free allocation)
#ifndef SB2_PLAN_MINB
#define SB2_PLAN_MINB 16
#endif
#ifndef SB2_EXEC_MINB
#define SB2_EXEC_MINB 8
#endif

__global__ void k_reset(const __grid_constant__ Ctx c) {
    if (threadIdx.x == 0) {
        c.counters[0] = 0;
        c.counters[7] = 0;
        c.counters[3] = 0;
        c.counters[8] = 0;
        c.counters[9] = 0;
        c.counters[11] = 0;
    }
}

__global__ void __launch_bounds__(BLOCK) k_pre(const __grid_constant__ Ctx c) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid < c.nTL) {
        tls_advance(c, tid);
    }
    // thread tid also handles lane lane_lo + tid and link link_lo + tid (the partition's visible range, sb2_types.h)
    const int i = tid < c.lane_hi - c.lane_lo ? c.lane_lo + tid : c.nL;
    // lane-ordered work list: block-wide exclusive scan of the lanes' vehicle counts, one atomic per block for the base.
    // The list has two segments: [0, counters[7]) filled forwards and [nV - counters[3], nV) filled from the end backwards.
    // Default split: lanes whose first link is red/yellow go to the second segment (measured on B200, 1M vehicles on the
    // traffic-light grid: k_plan 2.07 -> 1.44 ms, because a warp then holds vehicles that all stop at the red link after one
    // look-ahead iteration, or none).  -DSB2_SPLIT_INTERNAL (internal lanes in the second segment) measured no gain.
    __shared__ int warpSum[2][BLOCK / 32];
    __shared__ int blockBase[2];
#if defined(SB2_SPLIT_FIRST)
    __shared__ int warpSumF[2][BLOCK / 32];
    __shared__ int blockFirsts[2];
#endif
    const int nveh = (i < c.nL && lane_simulated(c, i)) ? c.lane_nveh[i] : 0;   // work list: lanes this rank simulates
#if defined(SB2_SPLIT_INTERNAL)
    const bool internal = i < c.nL && (c.lane_flags[i] & LF_INTERNAL) != 0;
#elif !defined(SB2_NO_SPLIT_RED)
    // second segment = lanes whose first link shows red or yellow: their vehicles mostly end planMove's look-ahead loop at that
    // link (stopAtRed), the others run 2-4 iterations.  (The state may be the one before or after this step's switch: the order
    // of the work list never changes a result.)
    bool internal = false;
    if (i < c.nL && c.lane_link_begin[i] < c.lane_link_begin[i + 1]) {
        const uint8_t s0 = c.link_state[c.lane_link_begin[i]];
        internal = state_red(s0) || state_yellow(s0);
    }
#else
    const bool internal = false;
#endif
#if defined(SB2_SPLIT_FIRST)   // experiment, off by default until measured on B200 (tools_dev/ab_first.sh)
    const int n0 = internal ? 0 : nveh, n1 = internal ? nveh : 0;
    // within the block's range of each segment the front vehicles of its lanes come first, their followers after them: a
    // warp then holds front vehicles (next-lane leaders, link leaders, foes) or followers (leader on the own lane), not both
    const int f0 = n0 > 0 ? 1 : 0, f1 = n1 > 0 ? 1 : 0;
    int incl0 = n0 - f0, incl1 = n1 - f1, inclF0 = f0, inclF1 = f1;
    for (int d = 1; d < 32; d <<= 1) {
        const int o0 = __shfl_up_sync(0xffffffffu, incl0, d);
        const int o1 = __shfl_up_sync(0xffffffffu, incl1, d);
        const int p0 = __shfl_up_sync(0xffffffffu, inclF0, d);
        const int p1 = __shfl_up_sync(0xffffffffu, inclF1, d);
        if ((threadIdx.x & 31) >= d) {
            incl0 += o0;
            incl1 += o1;
            inclF0 += p0;
            inclF1 += p1;
        }
    }
    if ((threadIdx.x & 31) == 31) {
        warpSum[0][threadIdx.x >> 5] = incl0;
        warpSum[1][threadIdx.x >> 5] = incl1;
        warpSumF[0][threadIdx.x >> 5] = inclF0;
        warpSumF[1][threadIdx.x >> 5] = inclF1;
    }
    __syncthreads();
    if (threadIdx.x < 2) {
        int tot = 0, totF = 0;
        for (int w = 0; w < BLOCK / 32; w++) {
            const int x = warpSum[threadIdx.x][w];
            warpSum[threadIdx.x][w] = tot;
            tot += x;
            const int y = warpSumF[threadIdx.x][w];
            warpSumF[threadIdx.x][w] = totF;
            totF += y;
        }
        blockFirsts[threadIdx.x] = totF;
        blockBase[threadIdx.x] = tot + totF > 0 ? atomicAdd(&c.counters[threadIdx.x == 0 ? 7 : 3], tot + totF) : 0;
    }
    __syncthreads();
    if (i < c.nL && nveh > 0) {
        if (!internal) {
            const int firstSlot = blockBase[0] + warpSumF[0][threadIdx.x >> 5] + inclF0 - f0;
            const int restSlot = blockBase[0] + blockFirsts[0] + warpSum[0][threadIdx.x >> 5] + incl0 - (n0 - f0);
            lane_summary(c, i, restSlot, 1, f0 ? firstSlot : -1);
        } else {
            const int firstSlot = c.nV - 1 - (blockBase[1] + warpSumF[1][threadIdx.x >> 5] + inclF1 - f1);
            const int restSlot = c.nV - 1 - (blockBase[1] + blockFirsts[1] + warpSum[1][threadIdx.x >> 5] + incl1 - (n1 - f1));
            lane_summary(c, i, restSlot, -1, f1 ? firstSlot : -1);
        }
    } else if (i < c.nL) {
        lane_summary(c, i, 0, 1);
    }
#else
    const int n0 = internal ? 0 : nveh, n1 = internal ? nveh : 0;
    int incl0 = n0, incl1 = n1;
    for (int d = 1; d < 32; d <<= 1) {
        const int o0 = __shfl_up_sync(0xffffffffu, incl0, d);
        const int o1 = __shfl_up_sync(0xffffffffu, incl1, d);
        if ((threadIdx.x & 31) >= d) {
            incl0 += o0;
            incl1 += o1;
        }
    }
    if ((threadIdx.x & 31) == 31) {
        warpSum[0][threadIdx.x >> 5] = incl0;
        warpSum[1][threadIdx.x >> 5] = incl1;
    }
    __syncthreads();
    if (threadIdx.x < 2) {
        int tot = 0;
        for (int w = 0; w < BLOCK / 32; w++) {
            const int x = warpSum[threadIdx.x][w];
            warpSum[threadIdx.x][w] = tot;
            tot += x;
        }
        blockBase[threadIdx.x] = tot > 0 ? atomicAdd(&c.counters[threadIdx.x == 0 ? 7 : 3], tot) : 0;
    }
    __syncthreads();
    if (i < c.nL) {
        if (!internal) {
            lane_summary(c, i, blockBase[0] + warpSum[0][threadIdx.x >> 5] + incl0 - n0, 1);
        } else {
            lane_summary(c, i, c.nV - 1 - (blockBase[1] + warpSum[1][threadIdx.x >> 5] + incl1 - n1), -1);
        }
    }
#endif
    if (tid < c.link_hi - c.link_lo) {
        c.link_app_head[c.cur][c.link_lo + tid] = -1;
    }
    if (tid < 3) {
        c.fleet[tid] = c.fleet_next[tid];
    }
}

__global__ void __launch_bounds__(BLOCK, SB2_PLAN_MINB) k_plan(const __grid_constant__ Ctx c) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < c.counters[7] || (i < c.nV && i >= c.nV - c.counters[3])) {
        veh_planMove(c, c.veh_order[c.cur][i], i);
    }
}

__global__ void __launch_bounds__(BLOCK, SB2_EXEC_MINB) k_execute(const __grid_constant__ Ctx c) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < c.counters[7] || (i < c.nV && i >= c.nV - c.counters[3])) {
        veh_executeMove_pass1(c, i);
    }
}

// second pass of executeMove for the (rare) vehicles whose right of way depends on the new speed of a foe the reference moves
// earlier in the step (MSLink::blockedAtTime under lastWasContRed, see sb2_link.h LIVE_PASS1): one CTA, rounds until all are done
__global__ void __launch_bounds__(256) k_execute2(const __grid_constant__ Ctx c) {
    const int n = c.counters[9];
    if (n == 0) {
        return;
    }
    __shared__ int left;
    for (int round = 0; round < 16; round++) {
        if (threadIdx.x == 0) {
            left = 0;
        }
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            if (!veh_execute_deferred(c, i)) {
                atomicAdd(&left, 1);
            }
        }
        __threadfence();
        __syncthreads();
        if (left == 0) {
            return;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        raise(c, ERR_DEFERRED_UNRESOLVED);
    }
}

__global__ void __launch_bounds__(BLOCK) k_remove(const __grid_constant__ Ctx c) {
    const int l = c.lane_lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (l < c.lane_hi && lane_simulated(c, l)) {
        lane_handoff_remove(c, l);
    }
}

__global__ void __launch_bounds__(BLOCK) k_integrate(const __grid_constant__ Ctx c) {
    const int l = c.lane_lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (l < c.lane_hi && lane_simulated(c, l)) {
        lane_handoff_integrate(c, l);
    }
}

// MSEdgeControl::detectCollisions(STAGE_MOVEMENTS): a collider is teleported (collision.action teleport, the default)
__global__ void __launch_bounds__(BLOCK) k_collide(const __grid_constant__ Ctx c) {
    const int l = c.lane_lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (l < c.lane_hi && lane_simulated(c, l)) {
        lane_detect_collisions(c, l);
    }
}

__global__ void __launch_bounds__(BLOCK) k_partials(const __grid_constant__ Ctx c) {
    const int l = c.lane_lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (l < c.lane_hi && lane_simulated(c, l)) {
        lane_sort_partials(c, l);
    }
}

// MSVehicleTransfer::checkInsertions: the teleporting vehicles are handled one after the other, in numerical id order
__global__ void k_transfer(const __grid_constant__ Ctx c) {
    if (blockIdx.x == 0 && threadIdx.x == 0 && c.counters[10] > 0) {
        transfer_checkInsertions(c);
    }
}

__global__ void __launch_bounds__(BLOCK) k_insert(const __grid_constant__ Ctx c) {
    const int l = c.lane_lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (l < c.lane_hi && c.dq_head[l] < c.dq_begin[l + 1] && lane_simulated(c, l)) {
        lane_insertVehicles(c, l);
    }
}

// running vehicles of this step -> counters[0] (overwritten) and the 64-bit vehicle-update accumulator
__global__ void __launch_bounds__(1024) k_count(const __grid_constant__ Ctx c, unsigned long long* updates) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    const bool on = v < c.nV && c.veh_lane[v] >= 0 && lane_auth(c, c.veh_lane[v]);
    // MSVehicleControl::getHaltingVehicleNo (MSVehicleControl.cpp:568-580): on the road and slower than SUMO_const_haltingSpeed
    const bool halting = on && c.veh_speed[v] < SB2_HALTING_SPEED;
    // one atomic per 1024-thread block and counter: with one per warp the 3 x 31 k same-address atomics of a 1 M-vehicle step
    // serialised in L2 and made this the slowest of the small kernels (59 us, profiles/r01_launches_summary.txt)
    const int nOn = __syncthreads_count(on);
    const int nHalting = __syncthreads_count(halting);
    if (threadIdx.x == 0 && nOn != 0) {
        atomicAdd(&c.counters[0], nOn);
        atomicAdd(updates, (unsigned long long)nOn);
        if (nHalting != 0) {
            atomicAdd(&c.counters[8], nHalting);
        }
    }
}

// ---- FCD snapshot (MSFCDExport::write): stream compaction of the on-road vehicle slots, in slot order ------------------
// pass 1: on-road vehicles per block
__global__ void __launch_bounds__(BLOCK) k_fcd_count(const __grid_constant__ Ctx c, int32_t* blockCount) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = __syncthreads_count(v < c.nV && c.veh_lane[v] >= 0 && lane_auth(c, c.veh_lane[v]));
    if (threadIdx.x == 0) {
        blockCount[blockIdx.x] = n;
    }
}
// pass 2: exclusive scan of the block counts by one CTA (chunks of 1024 with a running carry); total -> blockCount[nb]
__global__ void __launch_bounds__(1024) k_fcd_scan(int32_t* blockCount, int nb) {
    __shared__ int warpSum[32];
    __shared__ int carry;
    if (threadIdx.x == 0) {
        carry = 0;
    }
    __syncthreads();
    for (int base = 0; base < nb; base += 1024) {
        const int i = base + threadIdx.x;
        const int x = i < nb ? blockCount[i] : 0;
        int incl = x;
        for (int d = 1; d < 32; d <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, incl, d);
            if ((threadIdx.x & 31) >= d) {
                incl += o;
            }
        }
        if ((threadIdx.x & 31) == 31) {
            warpSum[threadIdx.x >> 5] = incl;
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            int w = warpSum[threadIdx.x];
            for (int d = 1; d < 32; d <<= 1) {
                const int o = __shfl_up_sync(0xffffffffu, w, d);
                if (threadIdx.x >= d) {
                    w += o;
                }
            }
            warpSum[threadIdx.x] = w;   // inclusive over warps
        }
        __syncthreads();
        const int warpBase = (threadIdx.x >> 5) > 0 ? warpSum[(threadIdx.x >> 5) - 1] : 0;
        if (i < nb) {
            blockCount[i] = carry + warpBase + incl - x;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            carry += warpSum[31];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        blockCount[nb] = carry;
    }
}
// pass 3: one row per on-road vehicle at its compacted position
__global__ void __launch_bounds__(BLOCK) k_fcd_write(const __grid_constant__ Ctx c, const int32_t* blockBase, sumob200_fcd_row* rows) {
    __shared__ int warpBase[BLOCK / 32];
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    const bool on = v < c.nV && c.veh_lane[v] >= 0 && lane_auth(c, c.veh_lane[v]);   // a partitioned run reports its own vehicles
    const unsigned mask = __ballot_sync(0xffffffffu, on);
    if ((threadIdx.x & 31) == 0) {
        warpBase[threadIdx.x >> 5] = __popc(mask);
    }
    __syncthreads();
    int base = blockBase[blockIdx.x];
    for (int w = 0; w < (int)(threadIdx.x >> 5); w++) {
        base += warpBase[w];
    }
    if (on) {
        const int at = base + __popc(mask & ((1u << (threadIdx.x & 31)) - 1u));
        const Pos2 p = veh_position(c, v);
        sumob200_fcd_row r;
        r.slot = c.veh_ext[v];
        r.lane = c.veh_lane[v];
        r.pos = c.veh_pos[v];
        r.speed = c.veh_speed[v];
        r.x = p.x;
        r.y = p.y;
        r.angle = naviDegree(veh_computeAngle(c, v, p));
        rows[at] = r;
    }
}

__global__ void __launch_bounds__(BLOCK) k_tripinfo(const __grid_constant__ Ctx c, sumob200_tripinfo* out) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < c.nV) {
        sumob200_tripinfo r;
        const int lane = c.veh_lane[v];
        const bool departed = lane != -1, arrived = lane == -2;
        r.depart_ms = departed ? c.veh_depart_time[v] : -1;
        r.depart_delay_ms = departed ? c.veh_depart_time[v] - c.veh_depart_param[v] : c.t - c.veh_depart_param[v];
        r.arrival_ms = arrived ? c.veh_arrival_time[v] : -1;
        r.duration_ms = departed ? (arrived ? c.veh_arrival_time[v] : c.t) - c.veh_depart_time[v] : 0;
        r.waiting_ms = c.veh_trip_wait[v];
        r.time_loss_ms = departed ? TIME2STEPS(c.veh_timeloss[v]) : 0;
        r.depart_lane = c.veh_departlane[v];
        r.arrival_lane = c.veh_arrival_lane[v];
        r.waiting_count = c.veh_trip_waitcount[v];
        r.teleport_arrived = (arrived && c.veh_trip_amwaiting[v] >= 2) ? c.veh_trip_amwaiting[v] - 1 : 0;
        r.depart_pos = c.veh_trip_departpos[v];
        r.depart_speed = c.veh_trip_departspeed[v];
        r.arrival_pos = c.veh_arrivalpos[v];
        r.arrival_speed = c.veh_speed[v];
        r.route_length = c.veh_trip_routelen[v] + (arrived ? c.veh_arrivalpos[v] : (departed ? c.veh_pos[v] : 0.));
        r.speed_factor = c.veh_speedfactor[v];
        out[c.veh_ext[v]] = r;
    }
}

// ---- summary aggregates (MSNet::writeSummaryOutput): per-block tree reduction in a fixed order, then one block over the block
// results -- deterministic, no floating-point atomics
__device__ __forceinline__ void summary_block_reduce(SummaryPart& p, SummaryPart* sh) {
    sh[threadIdx.x] = p;
    __syncthreads();
    for (int d = blockDim.x >> 1; d > 0; d >>= 1) {
        if ((int)threadIdx.x < d) {
            summary_add(sh[threadIdx.x], sh[threadIdx.x + d]);
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(BLOCK) k_summary(const __grid_constant__ Ctx c, SummaryPart* parts) {
    __shared__ SummaryPart sh[BLOCK];
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    SummaryPart p;
    summary_zero(p);
    if (v < c.nV) {
        summary_vehicle(c, v, c.t - c.DELTA_T, p);
    }
    summary_block_reduce(p, sh);
    if (threadIdx.x == 0) {
        parts[blockIdx.x] = sh[0];
    }
}

__global__ void __launch_bounds__(256) k_summary_final(const SummaryPart* parts, int n, SummaryPart* out) {
    __shared__ SummaryPart sh[256];
    SummaryPart p;
    summary_zero(p);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        summary_add(p, parts[i]);
    }
    summary_block_reduce(p, sh);
    if (threadIdx.x == 0) {
        *out = sh[0];
    }
}

// TraCI-style control between two steps (libsumo::TrafficLight::setPhase)
__global__ void k_tls_set_phase(const __grid_constant__ Ctx c, int tl, int phase) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        tls_set_phase(c, tl, phase);
    }
}

// libsumo::Vehicle::setSpeed for a batch of vehicle slots (speed < 0 removes the advice)
__global__ void k_set_speed(const __grid_constant__ Ctx c, const int32_t* slots, const double* speeds, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && slots[i] >= 0 && slots[i] < c.nV) {
        c.veh_traci_speed[c.ext2slot[slots[i]]] = speeds[i] >= 0. ? speeds[i] : -1.;
    }
}

__global__ void k_mirror(const __grid_constant__ Ctx c, sumob200_vehicle_state* out) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < c.nV) {
        sumob200_vehicle_state s;
        s.lane = c.veh_lane[v];
        s.route_pos = c.veh_routepos[v];
        s.pos = c.veh_pos[v];
        s.speed = c.veh_speed[v];
        out[c.veh_ext[v]] = s;
    }
}


// ---- halo exchange of the edge partition (sb2_halo.h)
__global__ void __launch_bounds__(BLOCK) k_halo_pack_lanes(const __grid_constant__ Ctx c, const HaloSide h, const uint8_t* imp_mask, void* msg) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < h.nExp) {
        halo_pack_lane(c, h, msg, i);
    }
    if (i == h.nExp) {   // one extra thread: teleporting vehicles that have moved into the receiver's strip
        halo_pack_transfers(c, h, imp_mask);
    }
}
__global__ void __launch_bounds__(BLOCK) k_halo_pack_vehicles(const __grid_constant__ Ctx c, const HaloSide h, void* msg) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    halo_pack_vehicle(c, h, msg, i);
}
__global__ void __launch_bounds__(BLOCK) k_halo_clear(const __grid_constant__ Ctx c, const HaloSide h) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < h.nImp) {
        halo_clear_lane(c, h, i);
    }
}
__global__ void __launch_bounds__(BLOCK) k_halo_unpack(const __grid_constant__ Ctx c, const HaloSide h, const void* msg) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < h.nImp) {
        halo_unpack_lane(c, msg, i);
    }
    halo_unpack_vehicle(c, msg, i);
}

// libsumo::Lane::getLastStep* of every lane (sb2_transfer.h:lane_stats)
__global__ void __launch_bounds__(BLOCK) k_lane_stats(const __grid_constant__ Ctx c, sumob200_lane_stats* out) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l < c.nL) {
        const LaneStats r = lane_stats(c, l);
        sumob200_lane_stats o;
        o.vehicle_number = r.vehicleNumber; o.halting_number = r.haltingNumber;
        o.mean_speed = r.meanSpeed; o.occupancy = r.occupancy; o.waiting_time = r.waitingTime;
        out[l] = o;
    }
}

// E1 detectors: the vehicles on a detector at the end of an output interval (sb2_e1.h)
__global__ void __launch_bounds__(BLOCK) k_e1_collect(const __grid_constant__ Ctx c) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < c.nV && c.nE1 > 0) {
        veh_e1_collect_on_detector(c, v);
    }
}

// E2 detectors: MSE2Collector::detectorUpdate after the step / writeXMLOutput + reset at the end of an interval (sb2_e2.h)
__global__ void __launch_bounds__(BLOCK) k_e2_update(const __grid_constant__ Ctx c) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d < c.nE2) {
        e2_detector_update(c, d);
    }
}
__global__ void __launch_bounds__(BLOCK) k_e2_readout(const __grid_constant__ Ctx c, int buf, E2Result* out) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d < c.nE2) {
        e2_detector_readout(c, d, buf, out);
    }
}

// ---- re-slotting (engine.cu reslot): renumber the on-road vehicles in lane order ------------------------------------------------
// per block of lanes: number of vehicles on them
__global__ void __launch_bounds__(BLOCK) k_rs_lanecount(const __grid_constant__ Ctx c, int32_t* blockCount) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    int n = l < c.nL ? c.lane_nveh[l] : 0;
    for (int d = 16; d > 0; d >>= 1) {
        n += __shfl_down_sync(0xffffffffu, n, d);
    }
    __shared__ int w[BLOCK / 32];
    if ((threadIdx.x & 31) == 0) {
        w[threadIdx.x >> 5] = n;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int i = 0; i < BLOCK / 32; i++) {
            t += w[i];
        }
        blockCount[blockIdx.x] = t;
    }
}
// order[j] = the j-th vehicle in (lane index, front -> back) order; blockBase = exclusive scan of k_rs_lanecount's counts
__global__ void __launch_bounds__(BLOCK) k_rs_laneorder(const __grid_constant__ Ctx c, const int32_t* blockBase, int32_t* order) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = l < c.nL ? c.lane_nveh[l] : 0;
    int incl = n;
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, incl, d);
        if ((threadIdx.x & 31) >= d) {
            incl += o;
        }
    }
    __shared__ int w[BLOCK / 32];
    if ((threadIdx.x & 31) == 31) {
        w[threadIdx.x >> 5] = incl;
    }
    __syncthreads();
    int base = blockBase[blockIdx.x] + incl - n;
    for (int i = 0; i < (int)(threadIdx.x >> 5); i++) {
        base += w[i];
    }
    if (l < c.nL) {
        for (int v = c.lane_first[l]; v >= 0; v = c.veh_behind[v]) {
            order[base++] = v;
        }
    }
}
// slots of the on-road vehicles, ascending (compaction; blockBase = exclusive scan of k_fcd_count-style counts)
__global__ void __launch_bounds__(BLOCK) k_rs_slotcount(const __grid_constant__ Ctx c, int32_t* blockCount) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = __syncthreads_count(v < c.nV && c.veh_lane[v] >= 0);
    if (threadIdx.x == 0) {
        blockCount[blockIdx.x] = n;
    }
}
__global__ void __launch_bounds__(BLOCK) k_rs_slotlist(const __grid_constant__ Ctx c, const int32_t* blockBase, int32_t* sorted) {
    __shared__ int warpBase[BLOCK / 32];
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    const bool on = v < c.nV && c.veh_lane[v] >= 0;
    const unsigned mask = __ballot_sync(0xffffffffu, on);
    if ((threadIdx.x & 31) == 0) {
        warpBase[threadIdx.x >> 5] = __popc(mask);
    }
    __syncthreads();
    int base = blockBase[blockIdx.x];
    for (int w = 0; w < (int)(threadIdx.x >> 5); w++) {
        base += warpBase[w];
    }
    if (on) {
        sorted[base + __popc(mask & ((1u << (threadIdx.x & 31)) - 1u))] = v;
    }
}
// the permutation: the j-th vehicle in lane order gets the j-th smallest on-road slot
__global__ void __launch_bounds__(BLOCK) k_rs_identity(int32_t* newOfOld, int32_t* oldOfNew, int n) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) {
        newOfOld[v] = v;
        oldOfNew[v] = v;
    }
}
__global__ void __launch_bounds__(BLOCK) k_rs_maps(const int32_t* order, const int32_t* sorted, const int32_t* count, int32_t* newOfOld, int32_t* oldOfNew) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < *count) {
        newOfOld[order[j]] = sorted[j];
        oldOfNew[sorted[j]] = order[j];
    }
}
// after the per-vehicle arrays have been permuted: every stored slot number follows
__global__ void __launch_bounds__(BLOCK) k_rs_remap(const __grid_constant__ Ctx c, const int32_t* newOfOld, int appBuf) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (size_t)c.nV) {
        const int a = c.veh_ahead[i], b = c.veh_behind[i];
        if (a >= 0) {
            c.veh_ahead[i] = newOfOld[a];
        }
        if (b >= 0) {
            c.veh_behind[i] = newOfOld[b];
        }
        c.ext2slot[c.veh_ext[i]] = (int32_t)i;
    }
    if (i < (size_t)c.nL) {
        if (c.lane_first[i] >= 0) {
            c.lane_first[i] = newOfOld[c.lane_first[i]];
            c.lane_last[i] = newOfOld[c.lane_last[i]];
        }
        for (int k = 0; k < c.lane_npartial[i] && k < MAX_PARTIAL; k++) {
            c.lane_partial[i * MAX_PARTIAL + k] = newOfOld[c.lane_partial[i * MAX_PARTIAL + k]];
        }
    }
    if (i < (size_t)c.nV * c.max_items) {
        // the approach table the next planMove reads (records of unused items hold stale but valid slot numbers)
        const int v = c.ap_rec[appBuf][i].veh;
        const int dec = v < 0 ? ~v : v;
        if (dec >= 0 && dec < c.nV) {
            const int nw = newOfOld[dec];
            c.ap_rec[appBuf][i].veh = v < 0 ? ~nw : nw;
        }
    }
}

inline int grid(int n) { return n > 0 ? (n + BLOCK - 1) / BLOCK : 1; }
inline const Ctx& ctx(const void* p) { return *static_cast<const Ctx*>(p); }   // same layout as the engine's sb2::Ctx (one header)

void l_reset(const void* c, cudaStream_t s) { k_reset<<<1, 32, 0, s>>>(ctx(c)); }
void l_pre(const void* c, int n, cudaStream_t s) { k_pre<<<grid(n < 3 ? 3 : n), BLOCK, 0, s>>>(ctx(c)); }
void l_plan(const void* c, int n, cudaStream_t s) { k_plan<<<grid(n), BLOCK, 0, s>>>(ctx(c)); }
void l_execute(const void* c, int n, cudaStream_t s) { k_execute<<<grid(n), BLOCK, 0, s>>>(ctx(c)); k_execute2<<<1, 256, 0, s>>>(ctx(c)); }
void l_remove(const void* c, int n, cudaStream_t s) { k_remove<<<grid(n), BLOCK, 0, s>>>(ctx(c)); }
void l_integrate(const void* c, int n, cudaStream_t s) { k_integrate<<<grid(n), BLOCK, 0, s>>>(ctx(c)); k_collide<<<grid(n), BLOCK, 0, s>>>(ctx(c)); k_partials<<<grid(n), BLOCK, 0, s>>>(ctx(c)); }
void l_insert(const void* c, int n, cudaStream_t s) { k_insert<<<grid(n), BLOCK, 0, s>>>(ctx(c)); k_transfer<<<1, 32, 0, s>>>(ctx(c)); }
void l_count(const void* c, int n, unsigned long long* updates, cudaStream_t s) { k_count<<<n > 0 ? (n + 1023) / 1024 : 1, 1024, 0, s>>>(ctx(c), updates); }
void l_mirror(const void* c, int n, sumob200_vehicle_state* out, cudaStream_t s) { k_mirror<<<grid(n), BLOCK, 0, s>>>(ctx(c), out); }
void l_tripinfo(const void* c, int n, sumob200_tripinfo* out, cudaStream_t s) { k_tripinfo<<<grid(n), BLOCK, 0, s>>>(ctx(c), out); }
void l_tls_set_phase(const void* c, int tl, int phase, cudaStream_t s) { k_tls_set_phase<<<1, 32, 0, s>>>(ctx(c), tl, phase); }
void l_set_speed(const void* c, const int32_t* slots, const double* speeds, int n, cudaStream_t s) { k_set_speed<<<grid(n), BLOCK, 0, s>>>(ctx(c), slots, speeds, n); }
int l_summary(const void* c, int n, void* parts, void* out, cudaStream_t s) {   // parts = NULL: number of bytes parts needs
    const int nb = grid(n);
    if (parts == nullptr) {
        return (int)sizeof(SummaryPart) * (nb + 1);
    }
    k_summary<<<nb, BLOCK, 0, s>>>(ctx(c), static_cast<SummaryPart*>(parts));
    k_summary_final<<<1, 256, 0, s>>>(static_cast<SummaryPart*>(parts), nb, static_cast<SummaryPart*>(out));
    return 2;
}
void l_fcd(const void* c, int n, int32_t* blocks, sumob200_fcd_row* rows, cudaStream_t s) {
    const int nb = grid(n);
    k_fcd_count<<<nb, BLOCK, 0, s>>>(ctx(c), blocks);
    k_fcd_scan<<<1, 1024, 0, s>>>(blocks, nb);
    k_fcd_write<<<nb, BLOCK, 0, s>>>(ctx(c), blocks, rows);
}

int l_halo_pack(const void* c, const void* side, const uint8_t* imp_mask, void* msg, cudaStream_t s) {
    const HaloSide& h = *static_cast<const HaloSide*>(side);
    cudaMemsetAsync(h.cursor, 0, sizeof(int32_t), s);
    k_halo_pack_lanes<<<grid(h.nExp + 1), BLOCK, 0, s>>>(ctx(c), h, imp_mask, msg);
    k_halo_pack_vehicles<<<grid(h.cap), BLOCK, 0, s>>>(ctx(c), h, msg);
    return 2;
}
int l_halo_unpack(const void* c, const void* side, const void* msg, cudaStream_t s) {
    const HaloSide& h = *static_cast<const HaloSide*>(side);
    k_halo_clear<<<grid(h.nImp), BLOCK, 0, s>>>(ctx(c), h);
    k_halo_unpack<<<grid(h.cap > h.nImp ? h.cap : h.nImp), BLOCK, 0, s>>>(ctx(c), h, msg);
    return 2;
}

// renumber the on-road vehicles in lane order: fills newOfOld / oldOfNew (scratch: blocks >= grid(nL) + grid(nV) + 2 ints, order / sorted nV ints)
int l_reslot_maps(const void* c, int nL, int nV, int32_t* blocksA, int32_t* blocksB, int32_t* order, int32_t* sorted, int32_t* newOfOld, int32_t* oldOfNew, cudaStream_t s) {
    const int nbL = grid(nL), nbV = grid(nV);
    k_rs_lanecount<<<nbL, BLOCK, 0, s>>>(ctx(c), blocksA);
    k_fcd_scan<<<1, 1024, 0, s>>>(blocksA, nbL);
    k_rs_laneorder<<<nbL, BLOCK, 0, s>>>(ctx(c), blocksA, order);
    k_rs_slotcount<<<nbV, BLOCK, 0, s>>>(ctx(c), blocksB);
    k_fcd_scan<<<1, 1024, 0, s>>>(blocksB, nbV);
    k_rs_slotlist<<<nbV, BLOCK, 0, s>>>(ctx(c), blocksB, sorted);
    k_rs_identity<<<nbV, BLOCK, 0, s>>>(newOfOld, oldOfNew, nV);
    k_rs_maps<<<nbV, BLOCK, 0, s>>>(order, sorted, blocksB + nbV, newOfOld, oldOfNew);
    return 8;
}
void l_e1_collect(const void* c, int n, cudaStream_t s) { k_e1_collect<<<grid(n), BLOCK, 0, s>>>(ctx(c)); }
void l_e2_update(const void* c, int n, cudaStream_t s) { k_e2_update<<<grid(n), BLOCK, 0, s>>>(ctx(c)); }
void l_e2_readout(const void* c, int n, int buf, void* out, cudaStream_t s) { k_e2_readout<<<grid(n), BLOCK, 0, s>>>(ctx(c), buf, static_cast<E2Result*>(out)); }
void l_lane_stats(const void* c, int n, sumob200_lane_stats* out, cudaStream_t s) { k_lane_stats<<<grid(n), BLOCK, 0, s>>>(ctx(c), out); }
int l_reslot_remap(const void* c, int nL, int nV, const int32_t* newOfOld, int appBuf, cudaStream_t s) {
    const size_t nDI = (size_t)nV * ctx(c).max_items;
    const size_t n = nDI > (size_t)nL ? nDI : (size_t)nL;
    k_rs_remap<<<(unsigned)((n + BLOCK - 1) / BLOCK), BLOCK, 0, s>>>(ctx(c), newOfOld, appBuf);
    return 1;
}

const sb2_kernelset kSet = {SB2_STR(SB2_KSET), BLOCK, sizeof(Ctx), l_reset, l_pre, l_plan, l_execute, l_remove, l_integrate, l_insert, l_count, l_mirror, l_fcd, l_tripinfo, l_summary, l_tls_set_phase, l_set_speed, l_halo_pack, l_halo_unpack, l_reslot_maps, l_reslot_remap, l_e1_collect, l_lane_stats, l_e2_update, l_e2_readout};

}  // namespace sb2_kset_<name>

extern "C" const sb2_kernelset* SB2_PASTE(sb2_kernelset_, SB2_KSET)(void) { return &SB2_PASTE(sb2_kset_, SB2_KSET)::kSet; }
