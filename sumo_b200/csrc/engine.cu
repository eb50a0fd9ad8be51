/* engine.cu -- libsumob200.so: the sm_100a kernels of the step pipeline and the C ABI of include/sumob200.h.
 *
 * Build: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -prec-div=true -prec-sqrt=true -lineinfo (see build.py).
 * -fmad=false keeps every fp64 expression in the reference's evaluation order without FMA contraction, so Krauss sigma=0
 * trajectories are bit-identical to the reference built with g++ -ffp-contract=off.
 *
 * One simulation step = 7 launches on one stream (no host synchronisation in between).  Every kernel takes the Ctx by value
 * as a __grid_constant__ parameter, so the ~150 array base pointers are read from the constant bank, not from HBM:
 *   k_pre       max(nL,nK,nTL) threads   TLS phase clock, per-lane summary (lengths in front, space till last standing),
 *                                        reset of this step's approach-table heads, fleet aggregate snapshot
 *   k_plan      nV threads               MSVehicle::planMove + approach registration           (sb2_plan2.h)
 *   (memset)    lane_npartial
 *   k_execute   nV threads               MSVehicle::executeMove, partial occupator registration (sb2_exec.h)
 *   k_remove    nL threads               pop vehicles that left each lane                       (sb2_lane.h)
 *   k_integrate nL threads               append arrivals, sort partial occupators, collisions
 *   k_insert    nL threads               MSInsertionControl::emitVehicles per departure lane    (sb2_insert.h)
 *   k_count     nV threads               running-vehicle count (summary / UPS numerator)
 * There is no CPU path in this library.
 */
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/sumob200.h"
#include "sb2_build.h"
#include "sb2_kernelset.h"
#include "sb2_fcdxml.h"
#include "sb2_summary.h"
#include "sb2_e1_host.h"
#include "sb2_halo.h"

using namespace sb2;

namespace {

thread_local std::string gLastError;

// dst[i] = src[oldOfNew[i]] for one per-vehicle array (re-slotting)
template<class T> __global__ void __launch_bounds__(256) k_permute(T* __restrict__ dst, const T* __restrict__ src, const int32_t* __restrict__ oldOfNew, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        dst[i] = src[oldOfNew[i]];
    }
}

#define CUDA_TRY(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) { \
        gLastError = std::string(#call) + ": " + cudaGetErrorString(_e); return SUMOB200_ECUDA; } } while (0)

}  // namespace

struct sumob200_engine {
    Model model;
    Ctx d;                                  // the Ctx passed by value to every kernel (pointers into device memory)
    const sb2_kernelset* ks = nullptr;      // kernel set chosen for the scenario (kernels.cu)
    std::vector<void*> dev;                 // device allocation of model.buffers[i]
    int device = 0;
    cudaStream_t stream = nullptr;
    unsigned long long* dUpdates = nullptr;
    sumob200_vehicle_state* dMirror = nullptr;
    void* dSummary = nullptr;   // SummaryPart scratch of sumob200_summary_now
    double* dReplay = nullptr;
    // FCD streaming: device rows + block counts, two pinned host buffers, a copy stream (D2H overlaps the next step)
    sumob200_fcd_row* dFcd[2] = {nullptr, nullptr};
    int32_t* dFcdBlocks[2] = {nullptr, nullptr};
    sumob200_fcd_row* hFcd[2] = {nullptr, nullptr};   // pinned
    int32_t* hFcdCount = nullptr;                     // [2], pinned
    int fcdCopiedRows[2] = {0, 0};                    // rows the asynchronous copy of each buffer covered
    int fcdLastCount = -1;                            // row count of the previous snapshot (sizes the next asynchronous copy)
    cudaStream_t copyStream = nullptr;
    cudaEvent_t fcdReady = nullptr, fcdCopied[2] = {nullptr, nullptr};
    int fcdBuf = 0, fcdPendingRows = -1;
    // edge partition (sb2_halo.h): per side 0 = towards lower columns, 1 = towards higher columns
    sb2::HaloSide halo[2];
    uint8_t* dLaneVis = nullptr;
    uint8_t* dImpMask[2] = {nullptr, nullptr};
    std::vector<void*> haloAlloc;
    bool partitioned = false;
    // re-slotting: every reslotPeriod steps the on-road vehicles are renumbered in lane order (0 = never)
    int reslotPeriod = 0;
    long long reslots = 0;
    int32_t* rsScratch = nullptr;   // blocksA, blocksB, order, sorted, newOfOld, oldOfNew
    bool debugSync = getenv("SUMOB200_DEBUG_SYNC") != nullptr;
    sumob200_lane_stats* dLaneStats = nullptr;
    sb2::E2Result* dE2Out = nullptr;
    std::vector<double> mdMinlenInit;   // SB2_DBL_MAX per lane: what md_minlen is reset to
    void* rsTmp = nullptr;          // one permuted array (16 B per vehicle)
    int64_t launches = 0;
    bool profiling = false;
    cudaEvent_t ev[6];
    double phaseMs[5] = {0, 0, 0, 0, 0};
    long long stepsDone = 0;
};

namespace {

int finishCreate(sumob200_engine* e, const FlatFile& ff, int device, sumob200_engine** out) {
    try {
        e->model.build(ff);
        // kernel set for this scenario (kernels.cu): code the fleet never executes is not compiled in
        const bool observers = e->model.h.nE1 > 0 || e->model.h.nE2 > 0 || e->model.h.md_on != 0;
        e->ks = e->model.anyIDM ? sb2_kernelset_full() : (observers ? sb2_kernelset_krauss() : sb2_kernelset_plain());
        if (e->ks->ctx_size != sizeof(Ctx)) {
            throw std::runtime_error("sumob200: kernel set was built with a different Ctx layout");
        }
    } catch (const std::invalid_argument& ex) {
        gLastError = ex.what();
        delete e;
        return SUMOB200_EINVAL;
    } catch (const std::exception& ex) {
        gLastError = ex.what();
        delete e;
        return SUMOB200_EUNSUPPORTED;
    }
    // (the scenario is validated first, so an unsupported feature is reported as such even where no GPU is present)
    int nDev = 0;
    if (cudaGetDeviceCount(&nDev) != cudaSuccess || nDev <= 0) {
        gLastError = "sumob200: no CUDA device available (this library has no CPU path)";
        delete e;
        return SUMOB200_ENODEVICE;
    }
    if (device < 0 || device >= nDev) {
        gLastError = "sumob200: invalid device index";
        delete e;
        return SUMOB200_EINVAL;
    }
    e->device = device;
    cudaError_t err = cudaSetDevice(device);
    if (err == cudaSuccess) {
        err = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
    }
    e->d = e->model.h;
    for (Buffer* b : e->model.buffers) {
        void* p = nullptr;
        if (err == cudaSuccess) {
            err = cudaMalloc(&p, b->host.size());
        }
        if (err == cudaSuccess) {
            err = cudaMemcpyAsync(p, b->host.data(), b->host.size(), cudaMemcpyHostToDevice, e->stream);
        }
        e->dev.push_back(p);
        *reinterpret_cast<void**>(reinterpret_cast<char*>(&e->d) + b->offset) = p;
    }
    e->d.cur = 1;   // the first step writes buffer 1 (the host emulation flips 0 -> 1 at the start of its first step)
    if (err == cudaSuccess) {
        err = cudaMalloc(&e->dUpdates, sizeof(unsigned long long));
    }
    if (err == cudaSuccess) {
        err = cudaMemsetAsync(e->dUpdates, 0, sizeof(unsigned long long), e->stream);
    }
    for (int i = 0; i < 6 && err == cudaSuccess; i++) {
        err = cudaEventCreate(&e->ev[i]);
    }
    if (err == cudaSuccess) {
        err = cudaStreamSynchronize(e->stream);
    }
    if (err != cudaSuccess) {
        gLastError = std::string("sumob200_create: ") + cudaGetErrorString(err);
        sumob200_destroy(e);
        return SUMOB200_ECUDA;
    }
    // the host copies of the big dynamic buffers are not needed any more
    for (Buffer* b : e->model.buffers) {
        if (b->dynamic) {
            std::vector<uint8_t>().swap(b->host);
        }
    }
    *out = e;
    return 0;
}

inline int gridOf(const sumob200_engine* e, int n) { return n > 0 ? (n + e->ks->block - 1) / e->ks->block : 1; }

// Re-slotting: renumber the on-road vehicles in (lane, front -> back) order so that vehicles that are neighbours on the road are
// neighbours in HBM (the per-vehicle kernels walk the lane-ordered work list: their state gathers become nearly coalesced).
// Every per-vehicle array -- static and dynamic alike, so no kernel needs an extra indirection -- is permuted, every stored slot
// number follows; veh_ext / ext2slot keep the external (load-order) index the C ABI speaks.  Nothing in the step depends on the
// slot number (ties are broken by numerical id / id rank, random draws are keyed by the external index): results are unchanged.
int reslot(sumob200_engine* e) {
    Ctx& c = e->d;
    const int nV = c.nV, nL = c.nL;
    const int nbL = gridOf(e, nL), nbV = gridOf(e, nV);
    cudaStream_t s = e->stream;
    if (e->rsScratch == nullptr) {
        CUDA_TRY(cudaMalloc(&e->rsScratch, sizeof(int32_t) * ((size_t)nbL + nbV + 4 + 4 * (size_t)(nV > 0 ? nV : 1))));
        CUDA_TRY(cudaMalloc(&e->rsTmp, 16 * (size_t)(nV > 0 ? nV : 1)));
    }
    int32_t* blocksA = e->rsScratch;
    int32_t* blocksB = blocksA + nbL + 2;
    int32_t* order = blocksB + nbV + 2;
    int32_t* sorted = order + nV;
    int32_t* newOfOld = sorted + nV;
    int32_t* oldOfNew = newOfOld + nV;
    e->launches += e->ks->reslot_maps(&e->d, nL, nV, blocksA, blocksB, order, sorted, newOfOld, oldOfNew, s);
    const int nb = (nV + 255) / 256;
    for (size_t i = 0; i < e->model.buffers.size(); i++) {
        const Buffer* b = e->model.buffers[i];
        if (b->name.compare(0, 4, "veh_") != 0 || b->name == "veh_order" || b->name == "veh_ndi") {
            continue;
        }
        char* base = static_cast<char*>(e->dev[i]);
        size_t rows = 1, bytes = b->elem;
        if (b->count == (size_t)nV) {
            // one element per vehicle
        } else if (b->name == "veh_further" && b->count == (size_t)nV * MAX_FURTHER) {
            bytes = b->elem * MAX_FURTHER;          // vehicle-major rows
        } else if (b->name == "veh_em" && b->count == (size_t)nV * 7) {
            rows = 7;                               // pollutant-major: seven arrays of nV
        } else if (b->name.compare(0, 7, "veh_e1_") == 0 && b->count == (size_t)nV * MAX_E1_REMINDERS) {
            rows = MAX_E1_REMINDERS;                // reminder-slot-major
        } else if (b->name.compare(0, 7, "veh_e2_") == 0 && b->count == (size_t)nV * MAX_E2_REMINDERS) {
            rows = MAX_E2_REMINDERS;
        } else {
            continue;                               // (veh_em of a scenario without emission devices: one dummy element)
        }
        for (size_t r = 0; r < rows; r++) {
            char* a = base + r * (size_t)nV * bytes;
            switch (bytes) {
                case 1: k_permute<uint8_t><<<nb, 256, 0, s>>>(static_cast<uint8_t*>(e->rsTmp), reinterpret_cast<const uint8_t*>(a), oldOfNew, nV); break;
                case 4: k_permute<int32_t><<<nb, 256, 0, s>>>(static_cast<int32_t*>(e->rsTmp), reinterpret_cast<const int32_t*>(a), oldOfNew, nV); break;
                case 8: k_permute<int64_t><<<nb, 256, 0, s>>>(static_cast<int64_t*>(e->rsTmp), reinterpret_cast<const int64_t*>(a), oldOfNew, nV); break;
                case 16: k_permute<int4><<<nb, 256, 0, s>>>(static_cast<int4*>(e->rsTmp), reinterpret_cast<const int4*>(a), oldOfNew, nV); break;
                default:
                    gLastError = "reslot: unexpected element size of " + b->name;
                    return SUMOB200_EINVAL;
            }
            CUDA_TRY(cudaMemcpyAsync(a, e->rsTmp, (size_t)nV * bytes, cudaMemcpyDeviceToDevice, s));
            e->launches += 1;
        }
    }
    // the approach table the next step reads is the one the last step wrote: buffer cur ^ 1 (cur was flipped at the end of the step)
    e->launches += e->ks->reslot_remap(&e->d, nL, nV, newOfOld, c.cur ^ 1, s);
    e->reslots++;
    CUDA_TRY(cudaGetLastError());
    return 0;
}

template<class T> int uploadArray(sumob200_engine* e, const T* host, size_t n, T** dev) {
    CUDA_TRY(cudaMalloc(dev, sizeof(T) * (n > 0 ? n : 1)));
    e->haloAlloc.push_back(*dev);
    if (n > 0) {
        CUDA_TRY(cudaMemcpy(*dev, host, sizeof(T) * n, cudaMemcpyHostToDevice));
    }
    return 0;
}

}  // namespace

extern "C" {

const char* sumob200_last_error(void) { return gLastError.c_str(); }
const char* sumob200_version(void) { return "sumob200 0.2 (sm_100a; Krauss + KraussOrig1 + IDM, single-lane edges, static TLS, teleports + collisions, HBEFA3/4, FCD streaming; kernel sets: full, krauss)"; }

int sumob200_create(const sumob200_array* arrays, int n_arrays, int device, sumob200_engine** out) {
    if (arrays == nullptr || out == nullptr || n_arrays <= 0) {
        gLastError = "sumob200_create: null argument";
        return SUMOB200_EINVAL;
    }
    FlatFile ff;
    for (int i = 0; i < n_arrays; i++) {
        const sumob200_array& a = arrays[i];
        if (a.name == nullptr || (a.count > 0 && a.data == nullptr) || a.elem_size == 0) {
            gLastError = "sumob200_create: malformed array descriptor";
            return SUMOB200_EINVAL;
        }
        FlatArray fa;
        fa.dtype = a.dtype;
        fa.elemSize = a.elem_size;
        fa.count = a.count;
        fa.bytes.resize(a.count * a.elem_size);
        if (!fa.bytes.empty()) {
            memcpy(fa.bytes.data(), a.data, fa.bytes.size());
        }
        ff.arrays[a.name] = std::move(fa);
    }
    return finishCreate(new sumob200_engine(), ff, device, out);
}

int sumob200_create_from_file(const char* flat_path, int device, sumob200_engine** out) {
    if (flat_path == nullptr || out == nullptr) {
        gLastError = "sumob200_create_from_file: null argument";
        return SUMOB200_EINVAL;
    }
    FlatFile ff;
    try {
        ff.load(flat_path);
    } catch (const std::exception& ex) {
        gLastError = ex.what();
        return SUMOB200_EINVAL;
    }
    return finishCreate(new sumob200_engine(), ff, device, out);
}

void sumob200_destroy(sumob200_engine* e) {
    if (e == nullptr) {
        return;
    }
    cudaSetDevice(e->device);
    for (void* p : e->dev) {
        if (p != nullptr) {
            cudaFree(p);
        }
    }
    if (e->dUpdates != nullptr) {
        cudaFree(e->dUpdates);
    }
    if (e->rsScratch != nullptr) {
        cudaFree(e->rsScratch);
        cudaFree(e->rsTmp);
        cudaFree(e->dLaneStats);
        cudaFree(e->dE2Out);
    }
    if (e->dSummary != nullptr) {
        cudaFree(e->dSummary);
    }
    if (e->dMirror != nullptr) {
        cudaFree(e->dMirror);
    }
    if (e->dReplay != nullptr) {
        cudaFree(e->dReplay);
    }
    if (e->dFcd[0] != nullptr) {
        cudaFree(e->dFcd[0]);
        cudaFree(e->dFcd[1]);
        cudaFree(e->dFcdBlocks[0]);
        cudaFree(e->dFcdBlocks[1]);
        cudaFreeHost(e->hFcd[0]);
        cudaFreeHost(e->hFcd[1]);
        cudaFreeHost(e->hFcdCount);
        cudaEventDestroy(e->fcdReady);
        cudaEventDestroy(e->fcdCopied[0]);
        cudaEventDestroy(e->fcdCopied[1]);
        cudaStreamDestroy(e->copyStream);
    }
    for (void* p : e->haloAlloc) {
        cudaFree(p);
    }
    if (e->stream != nullptr) {
        cudaStreamDestroy(e->stream);
    }
    delete e;
}

int sumob200_step(sumob200_engine* e, int n_steps, sumob200_step_result* out) {
    if (e == nullptr || n_steps < 0) {
        gLastError = "sumob200_step: bad argument";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    Ctx& c = e->d;
    const int nLv = c.lane_hi - c.lane_lo, nKv = c.link_hi - c.link_lo;   // visible lanes / links (everything unless partitioned)
    const int nPre = nLv > nKv ? (nLv > c.nTL ? nLv : c.nTL) : (nKv > c.nTL ? nKv : c.nTL);
    cudaStream_t s = e->stream;
    for (int i = 0; i < n_steps; i++) {
        const bool prof = e->profiling;
        if (prof) { cudaEventRecord(e->ev[0], s); }
        // SUMOB200_DEBUG_SYNC=1 (debugging only): wait after every launch group and name the one that failed
#define SB2_DBG(name) do { if (e->debugSync) { cudaError_t _d = cudaStreamSynchronize(s); if (_d != cudaSuccess) { \
        gLastError = std::string("step ") + std::to_string((long long)e->stepsDone) + ", after " name ": " + cudaGetErrorString(_d); return SUMOB200_ECUDA; } } } while (0)
        e->ks->reset(&e->d, s);   // counters[0] = running vehicles of this step, counters[7] = length of the work list
        e->ks->pre(&e->d, nPre, s);
        SB2_DBG("k_pre");
        if (prof) { cudaEventRecord(e->ev[1], s); }
        e->ks->plan(&e->d, c.nV, s);
        SB2_DBG("k_plan");
        if (prof) { cudaEventRecord(e->ev[2], s); }
        CUDA_TRY(cudaMemsetAsync(c.lane_npartial + c.lane_lo, 0, sizeof(int32_t) * (size_t)nLv, s));
        e->ks->execute(&e->d, c.nV, s);
        SB2_DBG("k_execute / k_execute2");
        if (prof) { cudaEventRecord(e->ev[3], s); }
        e->ks->remove(&e->d, nLv, s);
        SB2_DBG("k_remove");
        e->ks->integrate(&e->d, nLv, s);
        SB2_DBG("k_integrate / k_collide / k_partials");
        if (prof) { cudaEventRecord(e->ev[4], s); }
        e->ks->insert(&e->d, nLv, s);
        SB2_DBG("k_insert / k_transfer");
        e->ks->count(&e->d, c.nV, e->dUpdates, s);
        SB2_DBG("k_count");
        if (c.nE2 > 0) {
            e->ks->e2_update(&e->d, c.nE2, s);   // MSDetectorControl::updateDetectors -> MSE2Collector::detectorUpdate
            SB2_DBG("k_e2_update");
            e->launches += 1;
        }
        if (prof) {
            cudaEventRecord(e->ev[5], s);
            CUDA_TRY(cudaEventSynchronize(e->ev[5]));
            for (int p = 0; p < 5; p++) {
                float ms = 0;
                cudaEventElapsedTime(&ms, e->ev[p], e->ev[p + 1]);
                e->phaseMs[p] += ms;
            }
        }
        e->launches += 12;
        c.t += c.DELTA_T;
        c.cur ^= 1;
        e->stepsDone++;
        if (e->reslotPeriod > 0 && !e->partitioned && e->stepsDone % e->reslotPeriod == 0) {
            const int rc = reslot(e);
            if (rc != 0) {
                return rc;
            }
        }
    }
    CUDA_TRY(cudaGetLastError());
    if (out != nullptr) {
        int32_t counters[16];
        int32_t flags = 0;
        unsigned long long updates = 0;
        CUDA_TRY(cudaMemcpyAsync(counters, c.counters, sizeof(counters), cudaMemcpyDeviceToHost, s));
        CUDA_TRY(cudaMemcpyAsync(&flags, c.err_flags, sizeof(flags), cudaMemcpyDeviceToHost, s));
        CUDA_TRY(cudaMemcpyAsync(&updates, e->dUpdates, sizeof(updates), cudaMemcpyDeviceToHost, s));
        CUDA_TRY(cudaStreamSynchronize(s));
        out->time_ms = c.t;
        out->running = counters[0] + counters[10] + counters[11];
        out->teleporting = counters[10] + counters[11];
        out->discarded_total = counters[15];
        out->arrived_total = counters[1];
        out->departed_total = counters[2];
        out->collisions = counters[4];
        out->emergency_stops = counters[5];
        out->teleports = counters[6];
        out->flags = flags;
        out->halting = counters[8];
        out->vehicle_updates = (int64_t)updates;
    }
    return 0;
}

int sumob200_step_timed(sumob200_engine* e, int n_steps, sumob200_step_result* out, double* elapsed_ms) {
    if (e == nullptr || elapsed_ms == nullptr) {
        gLastError = "sumob200_step_timed: bad argument";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    cudaEvent_t a, b;
    CUDA_TRY(cudaEventCreate(&a));
    CUDA_TRY(cudaEventCreate(&b));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    CUDA_TRY(cudaEventRecord(a, e->stream));
    const int rc = sumob200_step(e, n_steps, nullptr);
    float ms = 0;
    cudaError_t err = rc == 0 ? cudaEventRecord(b, e->stream) : cudaSuccess;
    if (rc == 0 && err == cudaSuccess) {
        err = cudaEventSynchronize(b);
    }
    if (rc == 0 && err == cudaSuccess) {
        err = cudaEventElapsedTime(&ms, a, b);
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    if (rc != 0) {
        return rc;
    }
    if (err != cudaSuccess) {
        gLastError = std::string("sumob200_step_timed: ") + cudaGetErrorString(err);
        return SUMOB200_ECUDA;
    }
    *elapsed_ms = ms;
    return sumob200_step(e, 0, out);
}

int sumob200_mirror_all(sumob200_engine* e, sumob200_vehicle_state* out, int capacity) {
    if (e == nullptr || out == nullptr || capacity < e->d.nV) {
        gLastError = "sumob200_mirror_all: bad argument / capacity too small";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    const int nV = e->d.nV;
    if (e->dMirror == nullptr) {
        CUDA_TRY(cudaMalloc(&e->dMirror, sizeof(sumob200_vehicle_state) * (size_t)(nV > 0 ? nV : 1)));
    }
    e->ks->mirror(&e->d, nV, e->dMirror, e->stream);
    e->launches += 1;
    CUDA_TRY(cudaMemcpyAsync(out, e->dMirror, sizeof(sumob200_vehicle_state) * (size_t)nV, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    return 0;
}

int sumob200_summary_now(sumob200_engine* e, sumob200_summary* out) {
    if (e == nullptr || out == nullptr) {
        gLastError = "sumob200_summary_now: bad argument";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    const int nV = e->d.nV;
    const int bytes = e->ks->summary(&e->d, nV, nullptr, nullptr, e->stream);
    if (e->dSummary == nullptr) {
        CUDA_TRY(cudaMalloc(&e->dSummary, (size_t)bytes));
    }
    sb2::SummaryPart* res = reinterpret_cast<sb2::SummaryPart*>(static_cast<char*>(e->dSummary) + bytes) - 1;
    e->launches += e->ks->summary(&e->d, nV, e->dSummary, res, e->stream);
    sb2::SummaryPart p;
    int32_t counters[16];
    CUDA_TRY(cudaMemcpyAsync(&p, res, sizeof(p), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaMemcpyAsync(counters, e->d.counters, sizeof(counters), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    out->time_ms = e->d.t - e->d.DELTA_T;
    out->inserted = p.departed;
    out->running = p.count + p.teleporting;
    out->waiting = p.waiting;
    out->ended = p.ended;
    out->arrived = p.ended;
    out->collisions = counters[4];
    out->teleports = counters[6];
    out->halting = p.halting;
    out->stopped = 0;
    out->reserved = 0;
    out->mean_waiting_time = p.departed != 0 ? ((double)p.delay_ms / 1000.) / (double)p.departed : -1.;
    out->mean_travel_time = p.ended != 0 ? ((double)p.travel_ms / 1000.) / (double)p.ended : -1.;
    out->mean_speed = p.count > 0 ? p.speedSum / p.count : -1.;
    out->mean_speed_relative = p.count > 0 ? p.relSpeedSum / p.count : -1.;
    return 0;
}

int sumob200_tripinfo_all(sumob200_engine* e, sumob200_tripinfo* out, int capacity) {
    if (e == nullptr || out == nullptr || capacity < e->d.nV) {
        gLastError = "sumob200_tripinfo_all: bad argument / capacity too small";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    const int nV = e->d.nV;
    sumob200_tripinfo* d = nullptr;
    CUDA_TRY(cudaMalloc(&d, sizeof(sumob200_tripinfo) * (size_t)(nV > 0 ? nV : 1)));
    e->ks->tripinfo(&e->d, nV, d, e->stream);
    e->launches += 1;
    cudaError_t err = cudaMemcpyAsync(out, d, sizeof(sumob200_tripinfo) * (size_t)nV, cudaMemcpyDeviceToHost, e->stream);
    if (err == cudaSuccess) {
        err = cudaStreamSynchronize(e->stream);
    }
    cudaFree(d);
    if (err != cudaSuccess) {
        gLastError = std::string("sumob200_tripinfo_all: ") + cudaGetErrorString(err);
        return SUMOB200_ECUDA;
    }
    return 0;
}

int sumob200_emissions(sumob200_engine* e, double* out, int64_t capacity) {
    if (e == nullptr || out == nullptr || capacity < (int64_t)7 * e->d.nV) {
        gLastError = "sumob200_emissions: bad argument / capacity too small";
        return SUMOB200_EINVAL;
    }
    if (!e->model.anyEmissionDevice) {
        gLastError = "sumob200_emissions: no vehicle of this scenario carries the emission device";
        return SUMOB200_EUNSUPPORTED;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    CUDA_TRY(cudaMemcpyAsync(out, e->d.veh_em, sizeof(double) * 7 * (size_t)e->d.nV, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    if (e->reslots > 0) {
        // the device array is in memory (slot) order: hand it out in load order like everything else of this API
        const size_t nV = (size_t)e->d.nV;
        std::vector<int32_t> ext(nV);
        std::vector<double> tmp(nV);
        CUDA_TRY(cudaMemcpy(ext.data(), e->d.veh_ext, sizeof(int32_t) * nV, cudaMemcpyDeviceToHost));
        for (int p = 0; p < 7; p++) {
            double* row = out + (size_t)p * nV;
            for (size_t v = 0; v < nV; v++) {
                tmp[ext[v]] = row[v];
            }
            memcpy(row, tmp.data(), sizeof(double) * nV);
        }
    }
    return 0;
}

// ---- edge partition -------------------------------------------------------------------------------------------------------
int sumob200_set_partition(sumob200_engine* e, const uint8_t* lane_vis, const int32_t* const* exp_lanes, const int32_t* n_exp,
                           const int32_t* const* imp_lanes, const int32_t* n_imp, int capacity) {
    if (e == nullptr || lane_vis == nullptr || exp_lanes == nullptr || imp_lanes == nullptr || n_exp == nullptr || n_imp == nullptr || capacity <= 0 || e->partitioned) {
        gLastError = "sumob200_set_partition: bad argument (or called twice)";
        return SUMOB200_EINVAL;
    }
    const int nL = e->d.nL;
    for (int side = 0; side < 2; side++) {
        for (int i = 0; i < n_exp[side]; i++) {
            if (exp_lanes[side][i] < 0 || exp_lanes[side][i] >= nL || lane_vis[exp_lanes[side][i]] != LV_AUTH) {
                gLastError = "sumob200_set_partition: an exported lane is not one of the rank's own lanes";
                return SUMOB200_EINVAL;
            }
        }
        for (int i = 0; i < n_imp[side]; i++) {
            const int l = imp_lanes[side][i];
            if (l < 0 || l >= nL || (lane_vis[l] != LV_GHOST && lane_vis[l] != LV_HALO)) {
                gLastError = "sumob200_set_partition: an imported lane is not a ghost / halo lane";
                return SUMOB200_EINVAL;
            }
        }
    }
    CUDA_TRY(cudaSetDevice(e->device));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    if (uploadArray(e, lane_vis, (size_t)nL, &e->dLaneVis) != 0) {
        return SUMOB200_ECUDA;
    }
    e->d.lane_vis = e->dLaneVis;
    {
        int lo = nL, hi = 0;
        for (int l = 0; l < nL; l++) {
            if (lane_vis[l] != LV_NONE) {
                lo = l < lo ? l : lo;
                hi = l + 1;
            }
        }
        if (lo >= hi) {
            lo = 0;
            hi = 0;
        }
        e->d.lane_lo = lo;
        e->d.lane_hi = hi;
        e->d.link_lo = e->model.h.lane_link_begin[lo];
        e->d.link_hi = e->model.h.lane_link_begin[hi];
    }
    for (int side = 0; side < 2; side++) {
        sb2::HaloSide& h = e->halo[side];
        int32_t* dExp = nullptr;
        int32_t* dImp = nullptr;
        if (uploadArray(e, exp_lanes[side], (size_t)n_exp[side], &dExp) != 0 || uploadArray(e, imp_lanes[side], (size_t)n_imp[side], &dImp) != 0) {
            return SUMOB200_ECUDA;
        }
        std::vector<uint8_t> mask((size_t)nL, 0);
        for (int i = 0; i < n_imp[side]; i++) {
            mask[imp_lanes[side][i]] = 1;
        }
        if (uploadArray(e, mask.data(), (size_t)nL, &e->dImpMask[side]) != 0) {
            return SUMOB200_ECUDA;
        }
        h.exp_lanes = dExp; h.nExp = n_exp[side];
        h.imp_lanes = dImp; h.nImp = n_imp[side];
        h.cap = capacity;
        CUDA_TRY(cudaMalloc(&h.cursor, sizeof(int32_t)));
        e->haloAlloc.push_back(h.cursor);
        CUDA_TRY(cudaMalloc(&h.slots, sizeof(int32_t) * (size_t)capacity));
        e->haloAlloc.push_back(h.slots);
    }
    e->partitioned = true;
    return 0;
}

int64_t sumob200_halo_message_bytes(const sumob200_engine* e, int side, int outgoing) {
    if (e == nullptr || !e->partitioned || side < 0 || side > 1) {
        return 0;
    }
    const sb2::HaloSide& h = e->halo[side];
    return (int64_t)sb2::halo_message_bytes(outgoing ? h.nExp : h.nImp, h.cap);
}

int sumob200_halo_pack(sumob200_engine* e, int side, void* device_message, int max_records) {
    if (e == nullptr || !e->partitioned || side < 0 || side > 1 || device_message == nullptr) {
        gLastError = "sumob200_halo_pack: bad argument / no partition set";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    sb2::HaloSide h = e->halo[side];
    if (max_records > 0 && max_records < h.cap) {
        h.cap = max_records;   // the caller transports only this many records (both ends agree on the size from the previous step's count)
    }
    e->launches += e->ks->halo_pack(&e->d, &h, e->dImpMask[side], device_message, e->stream);
    if (e->debugSync) { CUDA_TRY(cudaStreamSynchronize(e->stream)); }
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int sumob200_halo_unpack(sumob200_engine* e, int side, const void* device_message) {
    if (e == nullptr || !e->partitioned || side < 0 || side > 1 || device_message == nullptr) {
        gLastError = "sumob200_halo_unpack: bad argument / no partition set";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    e->launches += e->ks->halo_unpack(&e->d, &e->halo[side], device_message, e->stream);
    if (e->debugSync) { CUDA_TRY(cudaStreamSynchronize(e->stream)); }
    CUDA_TRY(cudaGetLastError());
    return 0;
}

void* sumob200_stream(sumob200_engine* e) { return e == nullptr ? nullptr : (void*)e->stream; }
double* sumob200_fleet_aggregates(sumob200_engine* e) { return e == nullptr ? nullptr : e->d.fleet_next; }

int sumob200_num_vehicles(const sumob200_engine* e) { return e == nullptr ? 0 : e->d.nV; }
int sumob200_num_lanes(const sumob200_engine* e) { return e == nullptr ? 0 : e->d.nL; }
int64_t sumob200_time_ms(const sumob200_engine* e) { return e == nullptr ? 0 : e->d.t; }
int64_t sumob200_kernel_launches(const sumob200_engine* e) { return e == nullptr ? 0 : e->launches; }

int sumob200_vehicles_set_speed(sumob200_engine* e, const int32_t* slots, const double* speeds, int n) {
    if (e == nullptr || n < 0 || (n > 0 && (slots == nullptr || speeds == nullptr))) {
        gLastError = "sumob200_vehicles_set_speed: bad argument";
        return SUMOB200_EINVAL;
    }
    for (int i = 0; i < n; i++) {
        if (slots[i] < 0 || slots[i] >= e->d.nV) {
            gLastError = "sumob200_vehicles_set_speed: vehicle slot " + std::to_string(slots[i]) + " is not known";
            return SUMOB200_EINVAL;
        }
    }
    if (n == 0) {
        return 0;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    int32_t* dSlots = nullptr;
    double* dSpeeds = nullptr;
    CUDA_TRY(cudaMalloc(&dSlots, sizeof(int32_t) * (size_t)n));
    cudaError_t err = cudaMalloc(&dSpeeds, sizeof(double) * (size_t)n);
    if (err == cudaSuccess) {
        err = cudaMemcpyAsync(dSlots, slots, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, e->stream);
    }
    if (err == cudaSuccess) {
        err = cudaMemcpyAsync(dSpeeds, speeds, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, e->stream);
    }
    if (err == cudaSuccess) {
        e->d.traci_speed_used = 1;   // from now on k_plan / k_execute read veh_traci_speed
        e->ks->set_speed(&e->d, dSlots, dSpeeds, n, e->stream);
        e->launches += 1;
        err = cudaStreamSynchronize(e->stream);
    }
    cudaFree(dSlots);
    cudaFree(dSpeeds);
    if (err != cudaSuccess) {
        gLastError = std::string("sumob200_vehicles_set_speed: ") + cudaGetErrorString(err);
        return SUMOB200_ECUDA;
    }
    return 0;
}

int sumob200_num_tls(const sumob200_engine* e) { return e == nullptr ? 0 : e->d.nTL; }

int sumob200_tls_set_phase(sumob200_engine* e, int tls, int phase) {
    if (e == nullptr || tls < 0 || tls >= e->d.nTL) {
        gLastError = "sumob200_tls_set_phase: bad engine / traffic light index";
        return SUMOB200_EINVAL;
    }
    const int nPhases = e->model.h.tl_phase_begin[tls + 1] - e->model.h.tl_phase_begin[tls];
    if (phase < 0 || phase >= nPhases) {
        // the message of libsumo::TrafficLight::setPhase (TrafficLight.cpp:826)
        gLastError = "The phase index " + std::to_string(phase) + " is not in the allowed range [0," + std::to_string(nPhases - 1) + "].";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    e->ks->tls_set_phase(&e->d, tls, phase, e->stream);
    e->launches += 1;
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int sumob200_tls_get_phase(sumob200_engine* e, int tls, int* phase, int64_t* next_switch_ms) {
    if (e == nullptr || tls < 0 || tls >= e->d.nTL || phase == nullptr) {
        gLastError = "sumob200_tls_get_phase: bad argument";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    int32_t p = 0;
    int64_t ns = 0;
    CUDA_TRY(cudaMemcpyAsync(&p, e->d.tl_cur_phase + tls, sizeof(p), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaMemcpyAsync(&ns, e->d.tl_next_switch + tls, sizeof(ns), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    *phase = p;
    if (next_switch_ms != nullptr) {
        *next_switch_ms = ns;
    }
    return 0;
}

int sumob200_set_random_draws(sumob200_engine* e, const double* draws, int n) {
    if (e == nullptr || draws == nullptr || n != e->d.nV) {
        gLastError = "sumob200_set_random_draws: bad argument";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    if (e->dReplay == nullptr) {
        CUDA_TRY(cudaMalloc(&e->dReplay, sizeof(double) * (size_t)(n > 0 ? n : 1)));
    }
    CUDA_TRY(cudaMemcpyAsync(e->dReplay, draws, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    e->d.rng_replay = e->dReplay;
    e->d.rng_mode = 2;
    return 0;
}

int sumob200_set_random_mode(sumob200_engine* e, int mode) {
    if (e == nullptr || mode < 0 || mode > 1) {
        gLastError = "sumob200_set_random_mode: bad argument";
        return SUMOB200_EINVAL;
    }
    e->d.rng_mode = mode;
    return 0;
}

int sumob200_set_reslot_period(sumob200_engine* e, int steps) {
    if (e == nullptr || steps < 0) {
        gLastError = "sumob200_set_reslot_period: bad argument";
        return SUMOB200_EINVAL;
    }
    e->reslotPeriod = steps;
    return 0;
}

int sumob200_set_profiling(sumob200_engine* e, int enabled) {
    if (e == nullptr) {
        return SUMOB200_EINVAL;
    }
    e->profiling = enabled != 0;
    return 0;
}

int sumob200_phase_times(sumob200_engine* e, double* ms, int n_phases, int reset) {
    if (e == nullptr || ms == nullptr) {
        return SUMOB200_EINVAL;
    }
    for (int i = 0; i < n_phases && i < 5; i++) {
        ms[i] = e->phaseMs[i];
    }
    if (reset) {
        for (double& x : e->phaseMs) {
            x = 0;
        }
    }
    return 0;
}

int sumob200_fcd_format_xml(sumob200_engine* e, const sumob200_fcd_row* rows, int n, int64_t time_ms, char* buf, uint64_t capacity, uint64_t* length) {
    if (e == nullptr || (n > 0 && rows == nullptr) || length == nullptr || (capacity > 0 && buf == nullptr)) {
        gLastError = "sumob200_fcd_format_xml: bad argument";
        return SUMOB200_EINVAL;
    }
    const Model& m = e->model;
    if (m.veh_names.empty() || m.lane_names.empty() || m.vt_names.empty()) {
        gLastError = "sumob200_fcd_format_xml: the scenario was flattened without ids (veh_names / lane_names / vt_names)";
        return SUMOB200_EUNSUPPORTED;
    }
    for (int i = 0; i < n; i++) {
        if (rows[i].slot < 0 || rows[i].slot >= m.h.nV || rows[i].lane < 0 || rows[i].lane >= m.h.nL) {
            gLastError = "sumob200_fcd_format_xml: row with an invalid slot or lane";
            return SUMOB200_EINVAL;
        }
    }
    const FcdXmlNames nm = {&m.veh_names, &m.lane_names, &m.vt_names, m.h.veh_type, m.h.veh_idrank, fcdxml_precision(m.h.DELTA_T)};
    std::string out;
    fcdxml_timestep(nm, rows, n, time_ms, out);
    *length = out.size();
    if (capacity > 0) {
        memcpy(buf, out.data(), out.size() < capacity ? out.size() : capacity);
    }
    return 0;
}

int sumob200_fcd_begin(sumob200_engine* e) {
    if (e == nullptr) {
        gLastError = "sumob200_fcd_begin: null engine";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    const int nV = e->d.nV;
    const int nb = gridOf(e, nV);
    if (e->dFcd[0] == nullptr) {
        const size_t bytes = sizeof(sumob200_fcd_row) * (size_t)(nV > 0 ? nV : 1);
        for (int b = 0; b < 2; b++) {
            CUDA_TRY(cudaMalloc(&e->dFcd[b], bytes));
            CUDA_TRY(cudaMalloc(&e->dFcdBlocks[b], sizeof(int32_t) * (size_t)(nb + 1)));
            CUDA_TRY(cudaHostAlloc(&e->hFcd[b], bytes, cudaHostAllocDefault));
        }
        CUDA_TRY(cudaHostAlloc(&e->hFcdCount, 2 * sizeof(int32_t), cudaHostAllocDefault));
        CUDA_TRY(cudaStreamCreateWithFlags(&e->copyStream, cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreateWithFlags(&e->fcdReady, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&e->fcdCopied[0], cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&e->fcdCopied[1], cudaEventDisableTiming));
    }
    if (e->fcdPendingRows >= 0) {
        gLastError = "sumob200_fcd_begin: the previous snapshot was not collected with sumob200_fcd_end";
        return SUMOB200_EINVAL;
    }
    const int b = e->fcdBuf;
    cudaStream_t s = e->stream;
    // the D2H out of dFcd[b] two snapshots ago must be done before its rows are overwritten (it is, unless the host falls behind)
    CUDA_TRY(cudaStreamWaitEvent(s, e->fcdCopied[b], 0));
    e->ks->fcd(&e->d, nV, e->dFcdBlocks[b], e->dFcd[b], s);
    CUDA_TRY(cudaEventRecord(e->fcdReady, s));
    // rows + count travel on the copy stream (copy engine, no SM) while the caller launches the next step.  The number of rows is
    // only known on the device; waiting for it here would stall the pipeline (the next step could not be enqueued before this
    // snapshot was compacted).  So the copy is sized from the PREVIOUS snapshot's count plus head room -- the number of vehicles on
    // the road changes slowly -- and sumob200_fcd_end fetches the remainder in the rare case it was not enough.
    CUDA_TRY(cudaStreamWaitEvent(e->copyStream, e->fcdReady, 0));
    int guess = e->fcdLastCount < 0 ? nV : e->fcdLastCount + e->fcdLastCount / 16 + 1024;
    if (guess > nV) {
        guess = nV;
    }
    CUDA_TRY(cudaMemcpyAsync(e->hFcdCount + b, e->dFcdBlocks[b] + nb, sizeof(int32_t), cudaMemcpyDeviceToHost, e->copyStream));
    if (guess > 0) {
        CUDA_TRY(cudaMemcpyAsync(e->hFcd[b], e->dFcd[b], sizeof(sumob200_fcd_row) * (size_t)guess, cudaMemcpyDeviceToHost, e->copyStream));
    }
    e->fcdCopiedRows[b] = guess;
    e->launches += 3;
    CUDA_TRY(cudaEventRecord(e->fcdCopied[b], e->copyStream));
    e->fcdPendingRows = 0;
    return 0;
}

int sumob200_fcd_end(sumob200_engine* e, const sumob200_fcd_row** rows, int* n) {
    if (e == nullptr || rows == nullptr || n == nullptr || e->fcdPendingRows < 0) {
        gLastError = "sumob200_fcd_end: bad argument / no snapshot in flight";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    const int b = e->fcdBuf;
    CUDA_TRY(cudaEventSynchronize(e->fcdCopied[b]));
    const int count = e->hFcdCount[b];
    if (count > e->fcdCopiedRows[b]) {
        // more vehicles entered the road in one step than the head room covered: fetch the rest now
        CUDA_TRY(cudaMemcpyAsync(e->hFcd[b] + e->fcdCopiedRows[b], e->dFcd[b] + e->fcdCopiedRows[b],
                                 sizeof(sumob200_fcd_row) * (size_t)(count - e->fcdCopiedRows[b]), cudaMemcpyDeviceToHost, e->copyStream));
        CUDA_TRY(cudaStreamSynchronize(e->copyStream));
    }
    e->fcdLastCount = count;
    *rows = e->hFcd[b];
    *n = count;
    e->fcdPendingRows = -1;
    e->fcdBuf ^= 1;
    return 0;
}

int sumob200_fcd_snapshot(sumob200_engine* e, sumob200_fcd_row* rows, int capacity, int* n) {
    if (rows == nullptr || n == nullptr) {
        gLastError = "sumob200_fcd_snapshot: null argument";
        return SUMOB200_EINVAL;
    }
    int rc = sumob200_fcd_begin(e);
    if (rc != 0) {
        return rc;
    }
    const sumob200_fcd_row* src = nullptr;
    rc = sumob200_fcd_end(e, &src, n);
    if (rc != 0) {
        return rc;
    }
    memcpy(rows, src, sizeof(sumob200_fcd_row) * (size_t)(*n < capacity ? *n : capacity));
    return 0;
}

}  // extern "C"

// ---- E2 lane area detectors (sb2_e2.h) -----------------------------------------------------------------------------------
extern "C" int sumob200_e2_count(const sumob200_engine* e) { return e == nullptr ? 0 : e->d.nE2; }

extern "C" int sumob200_e2_interval(sumob200_engine* e, sumob200_e2_interval_t* out, int capacity) {
    static_assert(sizeof(sumob200_e2_interval_t) == sizeof(sb2::E2Result), "public and internal E2 record differ");
    if (e == nullptr || out == nullptr || capacity < e->d.nE2) {
        gLastError = "sumob200_e2_interval: bad argument / capacity too small";
        return SUMOB200_EINVAL;
    }
    const int nD = e->d.nE2;
    if (nD == 0) {
        return 0;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    if (e->dE2Out == nullptr) {
        CUDA_TRY(cudaMalloc(&e->dE2Out, sizeof(sb2::E2Result) * (size_t)nD));
    }
    // the notes of the last step that ran are in the buffer the step loop has flipped away from
    e->ks->e2_readout(&e->d, nD, e->d.cur ^ 1, e->dE2Out, e->stream);
    e->launches += 1;
    CUDA_TRY(cudaMemcpyAsync(out, e->dE2Out, sizeof(sb2::E2Result) * (size_t)nD, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    return 0;
}

extern "C" int sumob200_e2_notes(sumob200_engine* e, sumob200_e2_note* notes, int capacity, int* n, int32_t* counts) {
    if (e == nullptr || notes == nullptr || n == nullptr || capacity < 0) {
        gLastError = "sumob200_e2_notes: bad argument";
        return SUMOB200_EINVAL;
    }
    *n = 0;
    const int nD = e->d.nE2;
    if (nD == 0) {
        return 0;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    cudaStream_t s = e->stream;
    const int buf = e->d.cur ^ 1;   // the buffer of the last step that ran
    int32_t counters[16];
    std::vector<int32_t> head((size_t)nD), cnt((size_t)nD * sb2::E2C_N);
    CUDA_TRY(cudaMemcpyAsync(counters, e->d.counters, sizeof(counters), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaMemcpyAsync(head.data(), e->d.e2_head[buf], sizeof(int32_t) * (size_t)nD, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaMemcpyAsync(cnt.data(), e->d.e2_cnt, sizeof(int32_t) * cnt.size(), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    const int nN = counters[13 + buf] < e->d.e2_note_cap ? counters[13 + buf] : e->d.e2_note_cap;
    std::vector<sb2::E2Note> raw((size_t)nN);
    if (nN > 0) {
        CUDA_TRY(cudaMemcpyAsync(raw.data(), e->d.e2_notes[buf], sizeof(sb2::E2Note) * (size_t)nN, cudaMemcpyDeviceToHost, s));
        CUDA_TRY(cudaStreamSynchronize(s));
    }
    int k = 0;
    for (int d = 0; d < nD; d++) {
        for (int i = head[(size_t)d]; i >= 0 && i < nN; i = raw[(size_t)i].next) {
            if (k >= capacity) {
                gLastError = "sumob200_e2_notes: the notifications do not fit the caller's buffer";
                return SUMOB200_EINVAL;
            }
            const sb2::E2Note& r = raw[(size_t)i];
            sumob200_e2_note o;
            o.detector = d; o.vehicle = r.veh; o.dist_to_detector_end = r.dist; o.time_on_detector = r.timeOn;
            o.length_on_detector = r.lenOn; o.time_loss = r.timeLoss; o.speed = r.speed;
            notes[k++] = o;
        }
    }
    *n = k;
    if (counts != nullptr) {
        for (int d = 0; d < nD; d++) {
            for (int j = 0; j < 3; j++) {
                counts[3 * d + j] = cnt[(size_t)j * nD + d];
            }
        }
    }
    return 0;
}

// ---- E1 induction loops (sb2_e1.h) ---------------------------------------------------------------------------------------
extern "C" int sumob200_e1_count(const sumob200_engine* e) { return e == nullptr ? 0 : e->d.nE1; }

namespace {

// the event list of the running interval (+ pseudo events of the vehicles on a detector right now) and the entered counters,
// copied to the host; the device side starts a new interval (MSInductLoop::reset)
int e1Fetch(sumob200_engine* e, std::vector<sb2::E1Event>& ev, std::vector<int32_t>& entered) {
    const int nD = e->d.nE1;
    CUDA_TRY(cudaSetDevice(e->device));
    cudaStream_t s = e->stream;
    e->ks->e1_collect(&e->d, e->d.nV, s);
    e->launches += 1;
    int32_t counters[16];
    CUDA_TRY(cudaMemcpyAsync(counters, e->d.counters, sizeof(counters), cudaMemcpyDeviceToHost, s));
    entered.assign((size_t)nD, 0);
    CUDA_TRY(cudaMemcpyAsync(entered.data(), e->d.e1_entered, sizeof(int32_t) * (size_t)nD, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    const int n = counters[12] < e->d.e1_event_cap ? counters[12] : e->d.e1_event_cap;   // (an overflow has raised its flag)
    ev.resize((size_t)n);
    if (n > 0) {
        CUDA_TRY(cudaMemcpyAsync(ev.data(), e->d.e1_events, sizeof(sb2::E1Event) * (size_t)n, cudaMemcpyDeviceToHost, s));
    }
    CUDA_TRY(cudaMemsetAsync(e->d.counters + 12, 0, sizeof(int32_t), s));
    CUDA_TRY(cudaMemsetAsync(e->d.e1_entered, 0, sizeof(int32_t) * (size_t)nD, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    return 0;
}

}  // namespace

extern "C" int sumob200_e1_interval(sumob200_engine* e, int64_t begin_ms, int64_t end_ms, sumob200_e1_interval_t* out, int capacity) {
    static_assert(sizeof(sumob200_e1_interval_t) == sizeof(sb2::E1Result), "public and internal E1 record differ");
    if (e == nullptr || out == nullptr || capacity < e->d.nE1 || end_ms <= begin_ms) {
        gLastError = "sumob200_e1_interval: bad argument / capacity too small";
        return SUMOB200_EINVAL;
    }
    if (end_ms != e->d.t) {
        gLastError = "sumob200_e1_interval: the interval must end at the engine's current time";
        return SUMOB200_EINVAL;
    }
    if (e->d.nE1 == 0) {
        return 0;
    }
    std::vector<sb2::E1Event> ev;
    std::vector<int32_t> entered;
    const int rc = e1Fetch(e, ev, entered);
    if (rc != 0) {
        return rc;
    }
    sb2::e1_reduce(ev, entered.data(), e->d.nE1, e->model.h.e1_pos, e->model.h.e1_end, begin_ms, end_ms, reinterpret_cast<sb2::E1Result*>(out));
    return 0;
}

extern "C" int sumob200_e1_drain(sumob200_engine* e, sumob200_e1_event* events, int capacity, int* n, int32_t* entered_out) {
    static_assert(sizeof(sumob200_e1_event) == sizeof(sb2::E1Event), "public and internal E1 event differ");
    if (e == nullptr || events == nullptr || n == nullptr || capacity < 0) {
        gLastError = "sumob200_e1_drain: bad argument";
        return SUMOB200_EINVAL;
    }
    *n = 0;
    if (e->d.nE1 == 0) {
        return 0;
    }
    std::vector<sb2::E1Event> ev;
    std::vector<int32_t> entered;
    const int rc = e1Fetch(e, ev, entered);
    if (rc != 0) {
        return rc;
    }
    if ((int)ev.size() > capacity) {
        gLastError = "sumob200_e1_drain: " + std::to_string(ev.size()) + " events do not fit the caller's buffer (they are lost)";
        return SUMOB200_EINVAL;
    }
    sb2::e1_sort(ev);
    memcpy(events, ev.data(), sizeof(sb2::E1Event) * ev.size());
    *n = (int)ev.size();
    if (entered_out != nullptr) {
        memcpy(entered_out, entered.data(), sizeof(int32_t) * entered.size());
    }
    return 0;
}

// ---- libsumo::Lane::getLastStep* for all lanes ---------------------------------------------------------------------------------
extern "C" int sumob200_lane_stats_all(sumob200_engine* e, sumob200_lane_stats* out, int capacity) {
    if (e == nullptr || out == nullptr || capacity < e->d.nL) {
        gLastError = "sumob200_lane_stats_all: bad argument / capacity too small";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    const size_t nL = (size_t)e->d.nL;
    if (e->dLaneStats == nullptr) {
        CUDA_TRY(cudaMalloc(&e->dLaneStats, sizeof(sumob200_lane_stats) * (nL > 0 ? nL : 1)));
    }
    e->ks->lane_stats(&e->d, e->d.nL, e->dLaneStats, e->stream);
    e->launches += 1;
    CUDA_TRY(cudaMemcpyAsync(out, e->dLaneStats, sizeof(sumob200_lane_stats) * nL, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    return 0;
}

// ---- edgeData / laneData accumulators (sb2_meandata.h) -----------------------------------------------------------------
extern "C" int sumob200_meandata_interval(sumob200_engine* e, sumob200_lane_meandata* out, int capacity) {
    if (e == nullptr || out == nullptr || capacity < e->d.nL) {
        gLastError = "sumob200_meandata_interval: bad argument / capacity too small";
        return SUMOB200_EINVAL;
    }
    if (!e->d.md_on) {
        gLastError = "sumob200_meandata_interval: the scenario does not enable mean data (md_enabled)";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    const size_t nL = (size_t)e->d.nL;
    cudaStream_t s = e->stream;
    std::vector<double> sums(nL * MD_NSUMS), minlen(nL);
    std::vector<int32_t> counts(nL * MD_NCOUNTS);
    CUDA_TRY(cudaMemcpyAsync(sums.data(), e->d.md_sums, sizeof(double) * sums.size(), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaMemcpyAsync(minlen.data(), e->d.md_minlen, sizeof(double) * nL, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaMemcpyAsync(counts.data(), e->d.md_counts, sizeof(int32_t) * counts.size(), cudaMemcpyDeviceToHost, s));
    // MSLaneMeanDataValues::reset
    CUDA_TRY(cudaMemsetAsync(e->d.md_sums, 0, sizeof(double) * sums.size(), s));
    CUDA_TRY(cudaMemsetAsync(e->d.md_counts, 0, sizeof(int32_t) * counts.size(), s));
    e->mdMinlenInit.resize(nL, SB2_DBL_MAX);
    CUDA_TRY(cudaMemcpyAsync(e->d.md_minlen, e->mdMinlenInit.data(), sizeof(double) * nL, cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    for (size_t l = 0; l < nL; l++) {
        sumob200_lane_meandata& o = out[l];
        o.sample_seconds = sums[MD_SAMPLE * nL + l]; o.travelled_distance = sums[MD_DIST * nL + l];
        o.wait_seconds = sums[MD_WAIT * nL + l]; o.time_loss = sums[MD_TIMELOSS * nL + l];
        o.front_sample_seconds = sums[MD_FRONT_SAMPLE * nL + l]; o.front_travelled_distance = sums[MD_FRONT_DIST * nL + l];
        o.veh_length_sum = sums[MD_LENGTH_SUM * nL + l]; o.occupation_sum = sums[MD_OCCUPATION * nL + l];
        o.minimal_vehicle_length = minlen[l] == SB2_DBL_MAX ? -1. : minlen[l];
        o.departed = counts[MD_DEPARTED * nL + l]; o.arrived = counts[MD_ARRIVED * nL + l]; o.entered = counts[MD_ENTERED * nL + l];
        o.left = counts[MD_LEFT * nL + l]; o.teleported = counts[MD_TELEPORTED * nL + l]; o.reserved = 0;
    }
    return 0;
}

// ---- sumob200_summary_now_static: sumob200_summary_now without any allocation --------------------------------------------------
// No counterpart in the reference; added for callers that must not allocate device memory
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The per-block partial sums of k_summary and the
// reduced result live in ONE file-scope device array of a fixed size instead of the lazily cudaMalloc'ed e->dSummary: every GPU
// that loads the library holds its 917,560 bytes until exit, every call (of every engine on that GPU) shares it.  Calls that
// overlap on one GPU race on it; one engine's calls are ordered by its stream, different engines' calls must not run concurrently.
// Nothing in it has to start at zero: k_summary overwrites parts [0, nb), k_summary_final overwrites the result slot, and nothing
// else is read -- the result slot is cleared all the same, so that a launch that did not run cannot hand back the last call's sums.
namespace {

enum { SUMMARY_STATIC_MAX_PARTS = SUMOB200_SUMMARY_STATIC_MAX_VEHICLES / 128, SUMMARY_STATIC_MAX_DEVICES = 64 };
static __device__ __align__(256) unsigned char gSummaryScratch[sizeof(sb2::SummaryPart) * (SUMMARY_STATIC_MAX_PARTS + 1)];
// device address of gSummaryScratch per device (in host code the symbol's name is only a host-side stand-in, not a device address)
void* gSummaryScratchPtr[SUMMARY_STATIC_MAX_DEVICES];
std::mutex gSummaryScratchMutex;

// the array's address on `device` (which must be current); looked up once per device
int summaryScratchAddress(int device, void** out) {
    if (device < 0 || device >= SUMMARY_STATIC_MAX_DEVICES) {
        gLastError = "sumob200_summary_now_static: device index beyond " + std::to_string((int)SUMMARY_STATIC_MAX_DEVICES);
        return SUMOB200_EINVAL;
    }
    std::lock_guard<std::mutex> lock(gSummaryScratchMutex);
    if (gSummaryScratchPtr[device] == nullptr) {
        void* p = nullptr;
        const cudaError_t err = cudaGetSymbolAddress(&p, gSummaryScratch);
        if (err != cudaSuccess || p == nullptr) {
            gLastError = std::string("sumob200_summary_now_static: cudaGetSymbolAddress: ") + cudaGetErrorString(err);
            return SUMOB200_ECUDA;
        }
        gSummaryScratchPtr[device] = p;
    }
    *out = gSummaryScratchPtr[device];
    return 0;
}

}  // namespace

extern "C" int sumob200_summary_now_static(sumob200_engine* e, sumob200_summary* out) {
    if (e == nullptr || out == nullptr) {
        gLastError = "sumob200_summary_now_static: bad argument";
        return SUMOB200_EINVAL;
    }
    CUDA_TRY(cudaSetDevice(e->device));
    const int nV = e->d.nV;
    const int bytes = e->ks->summary(&e->d, nV, nullptr, nullptr, e->stream);   // (size query: launches nothing)
    if (nV > SUMOB200_SUMMARY_STATIC_MAX_VEHICLES || (size_t)bytes > sizeof(gSummaryScratch)) {
        gLastError = "sumob200_summary_now_static: " + std::to_string(nV) + " vehicle slots exceed the static scratch array's limit of "
                     + std::to_string((int)SUMOB200_SUMMARY_STATIC_MAX_VEHICLES) + " (use sumob200_summary_now)";
        return SUMOB200_EINVAL;
    }
    void* scratch = nullptr;
    const int rc = summaryScratchAddress(e->device, &scratch);
    if (rc != 0) {
        return rc;
    }
    sb2::SummaryPart* res = reinterpret_cast<sb2::SummaryPart*>(static_cast<char*>(scratch) + bytes) - 1;
    CUDA_TRY(cudaMemsetAsync(res, 0, sizeof(sb2::SummaryPart), e->stream));
    e->launches += e->ks->summary(&e->d, nV, scratch, res, e->stream);
    CUDA_TRY(cudaGetLastError());
    sb2::SummaryPart p;
    int32_t counters[16];
    CUDA_TRY(cudaMemcpyAsync(&p, res, sizeof(p), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaMemcpyAsync(counters, e->d.counters, sizeof(counters), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(cudaStreamSynchronize(e->stream));
    out->time_ms = e->d.t - e->d.DELTA_T;
    out->inserted = p.departed;
    out->running = p.count + p.teleporting;
    out->waiting = p.waiting;
    out->ended = p.ended;
    out->arrived = p.ended;
    out->collisions = counters[4];
    out->teleports = counters[6];
    out->halting = p.halting;
    out->stopped = 0;
    out->reserved = 0;
    out->mean_waiting_time = p.departed != 0 ? ((double)p.delay_ms / 1000.) / (double)p.departed : -1.;
    out->mean_travel_time = p.ended != 0 ? ((double)p.travel_ms / 1000.) / (double)p.ended : -1.;
    out->mean_speed = p.count > 0 ? p.speedSum / p.count : -1.;
    out->mean_speed_relative = p.count > 0 ? p.relSpeedSum / p.count : -1.;
    return 0;
}
