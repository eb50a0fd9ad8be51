/* sb2_build.h -- host side: turns the flattened scenario arrays (include/sumob200_flatfile.h) into the engine state.
 * Produces a list of named buffers plus a Ctx whose pointers refer to them; the CUDA engine uploads every buffer and
 * patches the pointers (engine.cu), the host emulation (tests/host_emul) uses them in place.
 *
 * Refuses, with a message, every option/feature the step kernels do not implement (SURVEY.md section 5 "config / flags").
 */
#pragma once
#include <cstddef>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
#include <algorithm>

#include "../../include/sumob200_flatfile.h"
#include "sb2_types.h"

namespace sb2 {

struct Buffer {
    std::string name;
    size_t offset;       // offset of the pointer inside Ctx
    std::vector<uint8_t> host;
    bool dynamic;        // written by kernels (state); static buffers are read-only after load
    size_t elem = 0, count = 0;   // element size and number of elements
};

class Model {
public:
    Ctx h;                         // host view
    bool anyIDM = false;           // some vType uses carFollowModel IDM (selects the kernel set)
    bool haveEmissionModel = false;
    bool anyEmissionDevice = false;
    std::vector<Buffer*> buffers;
    std::vector<std::string> lane_names, veh_names, vt_names;
    SUMOTime begin = 0, end = 0;

    Model() { memset(&h, 0, sizeof(h)); }
    ~Model() { for (Buffer* b : buffers) { delete b; } }
    Model(const Model&) = delete;
    Model& operator=(const Model&) = delete;

    template<class T> T* add(const char* name, size_t offset, size_t count, bool dynamic, T fill) {
        Buffer* b = new Buffer();
        b->name = name; b->offset = offset; b->dynamic = dynamic; b->elem = sizeof(T); b->count = count;
        b->host.resize(std::max<size_t>(count, 1) * sizeof(T));
        T* p = reinterpret_cast<T*>(b->host.data());
        for (size_t i = 0; i < std::max<size_t>(count, 1); i++) {
            p[i] = fill;
        }
        buffers.push_back(b);
        *reinterpret_cast<void**>(reinterpret_cast<char*>(&h) + offset) = p;
        return p;
    }
    template<class T> T* copy(const char* name, size_t offset, const FlatFile& ff, const char* flatName) {
        const FlatArray& a = ff.get(flatName);
        if (a.elemSize != sizeof(T)) {
            throw std::runtime_error(std::string("flat array '") + flatName + "' has unexpected element size");
        }
        T* p = add<T>(name, offset, a.count, false, T());
        if (a.count > 0) {
            memcpy(p, a.bytes.data(), a.count * sizeof(T));
        }
        return p;
    }

#define SB2_OFF(member) offsetof(Ctx, member)
#define SB2_COPY(T, member) copy<T>(#member, SB2_OFF(member), ff, #member)
#define SB2_COPY2(T, member, flatName) copy<T>(#member, SB2_OFF(member), ff, flatName)
#define SB2_DYN(T, member, count, fill) add<T>(#member, SB2_OFF(member), (count), true, (fill))
#define SB2_DYN2(T, member, idx, count, fill) add<T>(#member, SB2_OFF(member) + (idx) * sizeof(void*), (count), true, (fill))

    static std::vector<std::string> names(const FlatFile& ff, const std::string& prefix) {
        std::vector<std::string> out;
        const int32_t* b = ff.ptr<int32_t>(prefix + "_name_begin");
        const uint8_t* raw = ff.ptr<uint8_t>(prefix + "_names");
        const size_t n = ff.count(prefix + "_name_begin") - 1;
        for (size_t i = 0; i < n; i++) {
            out.emplace_back(reinterpret_cast<const char*>(raw) + b[i], reinterpret_cast<const char*>(raw) + b[i + 1]);
        }
        return out;
    }

    // ---- the arrays come from outside the library (C ABI): every index and every CSR offset array is range-checked before
    // anything is uploaded; a malformed array is std::invalid_argument (SUMOB200_EINVAL), never an out-of-bounds device access
    static void checkIndex(const FlatFile& ff, const char* name, int64_t lo, int64_t hi) {
        if (!ff.has(name)) {
            return;
        }
        const FlatArray& a = ff.get(name);
        if (a.elemSize != 4) {
            throw std::invalid_argument(std::string("flat array '") + name + "' must be int32");
        }
        const int32_t* p = a.as<int32_t>();
        for (uint64_t i = 0; i < a.count; i++) {
            if (p[i] < lo || p[i] >= hi) {
                throw std::invalid_argument(std::string("flat array '") + name + "': element " + std::to_string(i) + " = " + std::to_string(p[i]) +
                                            " is outside [" + std::to_string(lo) + ", " + std::to_string(hi) + ")");
            }
        }
    }
    static void checkCsr(const FlatFile& ff, const char* name, uint64_t rows, uint64_t payload) {
        const FlatArray& a = ff.get(name);
        if (a.elemSize != 4 || a.count != rows + 1) {
            throw std::invalid_argument(std::string("flat array '") + name + "' must have " + std::to_string(rows + 1) + " int32 offsets");
        }
        const int32_t* p = a.as<int32_t>();
        if (p[0] != 0 || (uint64_t)p[rows] != payload) {
            throw std::invalid_argument(std::string("flat array '") + name + "': offsets must run from 0 to " + std::to_string(payload));
        }
        for (uint64_t i = 0; i < rows; i++) {
            if (p[i] > p[i + 1]) {
                throw std::invalid_argument(std::string("flat array '") + name + "': offsets are not monotonic at " + std::to_string(i));
            }
        }
    }
    static void validate(const FlatFile& ff) {
        const int64_t nL = (int64_t)ff.count("lane_length"), nE = (int64_t)ff.count("edge_flags"), nK = (int64_t)ff.count("link_from");
        const int64_t nV = (int64_t)ff.count("veh_type"), nT = (int64_t)ff.count("vt_length"), nTL = (int64_t)ff.count("tl_cur_phase");
        const int64_t nR = (int64_t)ff.count("route_begin") - 1, nPh = (int64_t)ff.count("phase_dur");
        const int64_t nJ = 2 * (int64_t)nL + 2;   // junction indices are only compared with each other; an edge has two ends
        for (const char* per_lane : {"lane_speed", "lane_edge", "lane_flags", "lane_pred_logical", "lane_pred_normal", "lane_entry_link", "lane_to_junction", "lane_from_junction", "lane_geomfactor"}) {
            if ((int64_t)ff.count(per_lane) != nL) {
                throw std::invalid_argument(std::string("flat array '") + per_lane + "' must have one element per lane");
            }
        }
        for (const char* per_link : {"link_to", "link_via", "link_state", "link_dir", "link_length", "link_foevis", "link_radius", "link_flags", "link_index", "link_tl",
                                     "link_internal_before", "link_junction", "link_entry", "link_exit"}) {
            if ((int64_t)ff.count(per_link) != nK) {
                throw std::invalid_argument(std::string("flat array '") + per_link + "' must have one element per link");
            }
        }
        if ((int64_t)ff.count("link_response") != 4 * nK) {
            throw std::invalid_argument("flat array 'link_response' must have 4 words per link");
        }
        for (const char* per_veh : {"veh_route", "veh_numid", "veh_idrank", "veh_depart", "veh_speedfactor", "veh_arrivalpos", "veh_departpos", "veh_departpos_proc",
                                    "veh_departspeed", "veh_departspeed_proc", "veh_departlane", "veh_departlane_proc"}) {
            if ((int64_t)ff.count(per_veh) != nV) {
                throw std::invalid_argument(std::string("flat array '") + per_veh + "' must have one element per vehicle");
            }
        }
        checkCsr(ff, "edge_lane_begin", (uint64_t)nE, (uint64_t)nL);
        checkCsr(ff, "lane_link_begin", (uint64_t)nL, (uint64_t)nK);
        checkCsr(ff, "lane_inc_begin", (uint64_t)nL, ff.count("inc_lane"));
        checkCsr(ff, "lane_shape_begin", (uint64_t)nL, ff.count("shape_xy") / 2);
        checkCsr(ff, "link_foelink_begin", (uint64_t)nK, ff.count("foelink"));
        checkCsr(ff, "link_foelane_begin", (uint64_t)nK, ff.count("foelane"));
        checkCsr(ff, "route_begin", (uint64_t)nR, ff.count("route_edges"));
        checkCsr(ff, "tl_phase_begin", (uint64_t)nTL, (uint64_t)nPh);
        checkCsr(ff, "phase_state_begin", (uint64_t)nPh, ff.count("phase_states"));
        checkCsr(ff, "tl_link_begin", (uint64_t)nTL, ff.count("tllink_link"));
        if (ff.count("inc_link") != ff.count("inc_lane") || ff.count("inc_length") != ff.count("inc_lane") || ff.count("tllink_index") != ff.count("tllink_link")
                || ff.count("conf_lbc") != ff.count("foelane") || ff.count("conf_size") != ff.count("foelane") || ff.count("conf_flag") != ff.count("foelane")
                || ff.count("conf_foeidx") != ff.count("foelane")) {
            throw std::invalid_argument("flat arrays: payload arrays of one CSR structure differ in length");
        }
        checkIndex(ff, "lane_edge", 0, nE);
        checkIndex(ff, "lane_pred_logical", -1, nL); checkIndex(ff, "lane_pred_normal", -1, nL); checkIndex(ff, "lane_entry_link", -1, nK);
        checkIndex(ff, "lane_to_junction", -1, nJ); checkIndex(ff, "lane_from_junction", -1, nJ);
        checkIndex(ff, "inc_lane", 0, nL); checkIndex(ff, "inc_link", 0, nK);
        checkIndex(ff, "link_from", 0, nL); checkIndex(ff, "link_to", 0, nL); checkIndex(ff, "link_via", -1, nL);
        checkIndex(ff, "link_internal_before", -1, nL); checkIndex(ff, "link_entry", -1, nK); checkIndex(ff, "link_exit", -1, nK);
        checkIndex(ff, "link_tl", -1, nTL); checkIndex(ff, "link_index", -1, 256); checkIndex(ff, "link_junction", -1, nJ);
        checkIndex(ff, "foelink", 0, nK); checkIndex(ff, "foelane", -1, nL); checkIndex(ff, "conf_foeidx", -1, 1 << 20);
        checkIndex(ff, "tllink_link", 0, nK); checkIndex(ff, "tl_cur_phase", 0, nPh > 0 ? nPh : 1);
        checkIndex(ff, "route_edges", 0, nE);
        checkIndex(ff, "veh_type", 0, nT); checkIndex(ff, "veh_route", 0, nR);
        {
            // a traffic light's current phase and the state index of every controlled link must lie inside its own program
            const int32_t* pb = ff.ptr<int32_t>("tl_phase_begin");
            const int32_t* cur = ff.ptr<int32_t>("tl_cur_phase");
            const int32_t* sb = ff.ptr<int32_t>("phase_state_begin");
            const int32_t* lb = ff.ptr<int32_t>("tl_link_begin");
            const int32_t* li = ff.ptr<int32_t>("tllink_index");
            for (int64_t t = 0; t < nTL; t++) {
                if (pb[t + 1] > pb[t] && cur[t] >= pb[t + 1] - pb[t]) {
                    throw std::invalid_argument("flat array 'tl_cur_phase': phase index outside the program");
                }
                for (int ph = pb[t]; ph < pb[t + 1]; ph++) {
                    for (int i = lb[t]; i < lb[t + 1]; i++) {
                        if (li[i] < 0 || li[i] >= sb[ph + 1] - sb[ph]) {
                            throw std::invalid_argument("flat array 'tllink_index': state index outside the phase's state string");
                        }
                    }
                }
            }
        }
    }

    void build(const FlatFile& ff) {
        validate(ff);
        // ---------------- options
        h.DELTA_T = ff.scalar<int64_t>("opt_delta_t", 1000);
        h.TS = (double)h.DELTA_T / 1000.;
        begin = ff.scalar<int64_t>("opt_begin", 0);
        end = ff.scalar<int64_t>("opt_end", -1);
        h.t = begin;
        h.timeToImpatience = ff.scalar<int64_t>("opt_time_to_impatience", 180000);
        h.timeToTeleport = ff.scalar<int64_t>("opt_time_to_teleport", 300000);
        h.startupWaitThreshold = ff.scalar<int64_t>("opt_startup_wait_threshold", 2000);
        h.ignoreJunctionBlocker = ff.scalar<int64_t>("opt_ignore_junction_blocker", -1);
        h.maxDepartDelay = ff.scalar<int64_t>("opt_max_depart_delay", -1);
        h.extrapolateDepartPos = ff.scalar<int32_t>("opt_extrapolate_departpos", 0) != 0 ? 1 : 0;
        if (h.extrapolateDepartPos && (ff.has("e1_lane") || ff.has("e2_lane") || ff.scalar<int32_t>("md_enabled", 0) != 0)) {
            // (the reference notifies the move reminders of the fractional move: MSVehicle::executeFractionalMove -> workOnMoveReminders)
            throw std::runtime_error("sumob200: --extrapolate-departpos together with detectors / edgeData is not supported");
        }
        h.tlsYellowMinDecel = ff.scalar<double>("opt_tls_yellow_min_decel", 3.);
        if (ff.scalar<int32_t>("opt_sublane", 0) != 0) {
            throw std::runtime_error("sumob200: the sublane model (--lateral-resolution) is not supported");
        }
        if (ff.scalar<int32_t>("opt_semi_implicit_euler", 1) == 0) {
            throw std::runtime_error("sumob200: the ballistic update (--step-method.ballistic) is not supported");
        }
        if (ff.scalar<int32_t>("opt_using_internal_lanes", 1) == 0) {
            throw std::runtime_error("sumob200: --no-internal-links is not supported");
        }
        if (ff.scalar<double>("opt_lanechange_duration", 0.) > 0) {
            throw std::runtime_error("sumob200: --lanechange.duration > 0 is not supported");
        }
        h.nL = (int)ff.count("lane_length");
        h.nE = (int)ff.count("edge_flags");
        h.nK = (int)ff.count("link_from");
        h.nV = (int)ff.count("veh_type");
        h.nT = (int)ff.count("vt_length");
        h.nTL = (int)ff.count("tl_cur_phase");
        const int nL = h.nL, nK = h.nK, nV = h.nV, nT = h.nT;
        h.lane_lo = 0; h.lane_hi = nL; h.link_lo = 0; h.link_hi = nK;
        if (ff.has("lane_names") && ff.has("veh_names")) {   // ids are only needed for diagnostics and output writers
            lane_names = names(ff, "lane");
            veh_names = names(ff, "veh");
        }
        if (ff.has("vt_names")) {
            vt_names = names(ff, "vt");
        }
        // ---------------- lanes
        SB2_COPY(double, lane_length); SB2_COPY(double, lane_speed);
        SB2_COPY(int32_t, lane_edge); SB2_COPY(int32_t, lane_flags); SB2_COPY(int32_t, lane_link_begin);
        SB2_COPY(int32_t, lane_pred_logical); SB2_COPY(int32_t, lane_pred_normal); SB2_COPY(int32_t, lane_entry_link);
        SB2_COPY(int32_t, lane_inc_begin); SB2_COPY(int32_t, inc_lane); SB2_COPY(int32_t, inc_link); SB2_COPY(double, inc_length);
        SB2_COPY(int32_t, lane_to_junction); SB2_COPY(int32_t, lane_from_junction);
        SB2_COPY(int32_t, edge_lane_begin);
        SB2_COPY(int32_t, lane_shape_begin); SB2_COPY(double, shape_xy); SB2_COPY(double, lane_geomfactor);
        {
            // order keys of MSEdgeControl's lane containers: numerical id (myChangedStateLanes) and id string order (myWithVehicles2Integrate)
            int32_t* numid = add<int32_t>("lane_numid", SB2_OFF(lane_numid), nL, false, 0);
            int32_t* rank = add<int32_t>("lane_idrank", SB2_OFF(lane_idrank), nL, false, 0);
            for (int l = 0; l < nL; l++) {
                numid[l] = ff.has("lane_numid") ? ff.ptr<int32_t>("lane_numid")[l] : l;
                rank[l] = l;
            }
            if (ff.has("lane_idrank")) {
                memcpy(rank, ff.ptr<int32_t>("lane_idrank"), sizeof(int32_t) * (size_t)nL);
            } else if ((int)lane_names.size() == nL) {
                std::vector<int32_t> byName(nL);
                for (int l = 0; l < nL; l++) {
                    byName[l] = l;
                }
                std::sort(byName.begin(), byName.end(), [&](int a, int b) { return lane_names[a] < lane_names[b]; });
                for (int i = 0; i < nL; i++) {
                    rank[byName[i]] = i;
                }
            }
        }
        SB2_DYN(int64_t, lane_actkey, nL, (int64_t)0); SB2_DYN(uint8_t, lane_changed, nL, (uint8_t)0);
        SB2_DYN(int32_t, lane_first, nL, -1); SB2_DYN(int32_t, lane_last, nL, -1); SB2_DYN(int32_t, lane_nveh, nL, 0);
        SB2_DYN(double, lane_brutto, nL, 0.);
        SB2_DYN(int32_t, lane_npartial, nL, 0); SB2_DYN(int32_t, lane_partial, (size_t)nL * MAX_PARTIAL, -1);
        SB2_DYN(int32_t, lane_incoming, nL, -1);
        SB2_DYN(int32_t, ls_lastveh, nL, -1); SB2_DYN(double, ls_space, nL, 0.); SB2_DYN(uint8_t, ls_found_stopped, nL, (uint8_t)0);
        for (int e = 0; e < h.nE; e++) {
            const int nl = h.edge_lane_begin[e + 1] - h.edge_lane_begin[e];
            const bool internal = (ff.ptr<int32_t>("edge_flags")[e] & LF_INTERNAL) != 0;
            if (nl > 1 && !internal) {
                throw std::runtime_error("sumob200: multi-lane edges (lane changing, LC2013) are not implemented yet");
            }
        }
        // ---------------- links
        SB2_COPY(int32_t, link_from); SB2_COPY(int32_t, link_to); SB2_COPY(int32_t, link_via);
        {
            const FlatArray& a = ff.get("link_state");
            uint8_t* p = SB2_DYN(uint8_t, link_state, a.count, (uint8_t)0);
            if (a.count > 0) {
                memcpy(p, a.bytes.data(), a.count);
            }
            for (size_t k = 0; k < a.count; k++) {
                if (p[k] == 'Z') {
                    throw std::runtime_error("sumob200: zipper links are not supported");
                }
                // both junction types draw from the lane RNG streams inside planMove / setApproaching (right-before-left deadlock
                // breaker MSVehicle.cpp:3002-3020, all-way-stop arrival tie breaker MSVehicle.cpp:5500-5502): not reproducible
                // without the reference's stream order, so the scenario is refused instead of diverging silently
                if (p[k] == 'w') {
                    throw std::runtime_error("sumob200: allway_stop junctions are not supported (their arrival-time tie breaker draws from the lane RNG)");
                }
                if (p[k] == '=') {
                    throw std::runtime_error("sumob200: right_before_left junctions are not supported (their deadlock breaker draws from the lane RNG)");
                }
            }
        }
        SB2_COPY(uint8_t, link_dir);
        SB2_COPY(double, link_length); SB2_COPY(double, link_foevis); SB2_COPY(double, link_radius);
        SB2_COPY(int32_t, link_flags); SB2_COPY(int32_t, link_index); SB2_COPY(int32_t, link_tl);
        SB2_COPY(int32_t, link_internal_before); SB2_COPY(int32_t, link_junction); SB2_COPY(int32_t, link_entry); SB2_COPY(int32_t, link_exit);
        SB2_COPY(int32_t, link_foelink_begin); SB2_COPY(int32_t, foelink);
        SB2_COPY(int32_t, link_foelane_begin); SB2_COPY(int32_t, foelane);
        SB2_COPY(double, conf_lbc); SB2_COPY(double, conf_size); SB2_COPY(int32_t, conf_flag); SB2_COPY(int32_t, conf_foeidx);
        SB2_COPY(uint64_t, link_response);
        {
            // MSLink::lastWasContMajor / lastWasContState (MSLink.cpp:1353-1397): for a link whose via lane's logical
            // predecessor `pred` is internal, the link pred2 -> pred (pred2 = logical predecessor of pred)
            int32_t* cp = add<int32_t>("link_contpred", SB2_OFF(link_contpred), nK, false, -1);
            for (int k = 0; k < nK; k++) {
                const int via = h.link_via[k];
                if (via < 0) {
                    continue;
                }
                const int pred = h.lane_pred_logical[via];
                if (pred < 0 || !(h.lane_flags[pred] & LF_INTERNAL)) {
                    continue;
                }
                const int pred2 = h.lane_pred_logical[pred];
                if (pred2 < 0) {
                    continue;
                }
                for (int q = h.lane_link_begin[pred2]; q < h.lane_link_begin[pred2 + 1]; q++) {
                    // MSLane::getLinkTo(pred): pred is internal -> match the via lane
                    if (h.link_via[q] == pred) {
                        cp[k] = q;
                        break;
                    }
                }
            }
        }
        SB2_DYN(int64_t, link_last_change, nK, (int64_t)(INT64_MIN / 2));
        {
            uint8_t* lg = SB2_DYN(uint8_t, link_last_green, nK, (uint8_t)'g');
            (void)lg;
        }
        SB2_DYN2(int32_t, link_app_head, 0, nK, -1); SB2_DYN2(int32_t, link_app_head, 1, nK, -1);
        // ---------------- traffic lights
        SB2_COPY(int32_t, tl_phase_begin); SB2_COPY2(int64_t, phase_dur, "phase_dur"); SB2_COPY(int32_t, phase_state_begin); SB2_COPY(uint8_t, phase_states);
        {
            const FlatArray& a = ff.get("tl_cur_phase");
            int32_t* p = SB2_DYN(int32_t, tl_cur_phase, a.count, 0);
            if (a.count > 0) {
                memcpy(p, a.bytes.data(), a.count * 4);
            }
            const FlatArray& b = ff.get("tl_next_switch");
            int64_t* q = SB2_DYN(int64_t, tl_next_switch, b.count, (int64_t)0);
            if (b.count > 0) {
                memcpy(q, b.bytes.data(), b.count * 8);
            }
            if (ff.has("tl_static")) {
                const int32_t* st = ff.ptr<int32_t>("tl_static");
                for (size_t i = 0; i < ff.count("tl_static"); i++) {
                    if (!st[i]) {
                        throw std::runtime_error("sumob200: only static traffic light programs are supported");
                    }
                }
            }
        }
        SB2_COPY(int32_t, tl_link_begin); SB2_COPY(int32_t, tllink_index); SB2_COPY(int32_t, tllink_link);
        // ---------------- vTypes
        SB2_COPY(double, vt_length); SB2_COPY(double, vt_mingap); SB2_COPY(double, vt_maxspeed); SB2_COPY(double, vt_desiredmaxspeed);
        SB2_COPY(double, vt_accel); SB2_COPY(double, vt_decel); SB2_COPY(double, vt_emergencydecel); SB2_COPY(double, vt_apparentdecel);
        SB2_COPY(double, vt_tau); SB2_COPY(double, vt_sigma); SB2_COPY(double, vt_impatience);
        SB2_COPY2(int64_t, vt_startupdelay, "vt_startupdelay");
        SB2_COPY(double, vt_idm_delta); SB2_COPY(double, vt_idm_stepping);
        {
            const int32_t* tag = ff.ptr<int32_t>("vt_cfmodel");
            int32_t* m = add<int32_t>("vt_cfmodel", SB2_OFF(vt_cfmodel), nT, false, 0);
            const int tagKrauss = ff.scalar<int32_t>("opt_tag_cf_krauss", -1);
            const int tagOrig = ff.scalar<int32_t>("opt_tag_cf_krauss_orig1", -1);
            const int tagIDM = ff.scalar<int32_t>("opt_tag_cf_idm", -1);
            for (int i = 0; i < nT; i++) {
                if (tag[i] == tagKrauss) {
                    m[i] = CF_KRAUSS;
                } else if (tag[i] == tagOrig) {
                    m[i] = CF_KRAUSS_ORIG1;
                    anyIDM = true;   // (selects the `full` kernel set: the `krauss` set is compiled without the other models)
                } else if (tag[i] == tagIDM) {
                    m[i] = CF_IDM;
                    anyIDM = true;
                } else {
                    throw std::runtime_error("sumob200: unsupported car-following model (supported: Krauss, KraussOrig1, IDM)");
                }
            }
            SB2_COPY(double, vt_collision_mingap_factor);
            // emission model (optional arrays: scenarios flattened without them carry no emission device)
            haveEmissionModel = ff.has("vt_em_model") && ff.has("vt_em_coef");
            if (haveEmissionModel) {
                SB2_COPY(int32_t, vt_em_model); SB2_COPY(double, vt_em_coef); SB2_COPY(double, vt_em_scale);
                SB2_COPY(double, vt_em_mass); SB2_COPY(double, vt_em_incl);
            } else {
                SB2_DYN(int32_t, vt_em_model, nT, 0); SB2_DYN(double, vt_em_coef, (size_t)nT * 49, 0.); SB2_DYN(double, vt_em_scale, (size_t)nT * 7, 1.);
                SB2_DYN(double, vt_em_mass, nT, 0.); SB2_DYN(double, vt_em_incl, nT, 0.);
            }
            bool anySigma = false;
            for (int i = 0; i < nT; i++) {
                anySigma = anySigma || h.vt_sigma[i] != 0.;
            }
            h.rng_mode = anySigma ? 1 : 0;
            h.rng_seed = (uint64_t)ff.scalar<int64_t>("opt_seed", 23423);
            h.rng_replay = nullptr;
            h.traci_speed_used = 0;
            SB2_DYN(double, veh_traci_speed, nV, -1.);
        }
        // ---------------- routes, vehicles
        SB2_COPY(int32_t, route_begin); SB2_COPY(int32_t, route_edges);
        SB2_COPY(int32_t, veh_type); SB2_COPY(int32_t, veh_route); SB2_COPY(int32_t, veh_numid); SB2_COPY(int32_t, veh_idrank);
        SB2_COPY(double, veh_speedfactor); SB2_COPY(double, veh_arrivalpos);
        if (ff.has("veh_arrivalspeed")) {
            if ((int)ff.count("veh_arrivalspeed") != nV) {
                throw std::invalid_argument("sumob200: veh_arrivalspeed must have one element per vehicle");
            }
            SB2_COPY(double, veh_arrivalspeed);
        } else {
            add<double>("veh_arrivalspeed", SB2_OFF(veh_arrivalspeed), nV, false, -1.);
        }
        SB2_COPY(double, veh_departpos); SB2_COPY(int32_t, veh_departpos_proc);
        {
            // depart time = MSVehicleContainer bucket (rounded up to the step grid, MSVehicleContainer.cpp:80-81)
            const int64_t* d = ff.ptr<int64_t>("veh_depart");
            int64_t* o = add<int64_t>("veh_depart", SB2_OFF(veh_depart), nV, false, (int64_t)0);
            SB2_COPY2(SUMOTime, veh_depart_param, "veh_depart");
            for (int v = 0; v < nV; v++) {
                const int64_t delay = d[v] % h.DELTA_T;
                o[v] = d[v] + (delay == 0 ? 0 : h.DELTA_T - delay);
            }
            // depart speed: resolve DEFAULT (MSLane::getDepartSpeed, MSLane.cpp:609-614)
            const int32_t* sp = ff.ptr<int32_t>("veh_departspeed_proc");
            const double* sv = ff.ptr<double>("veh_departspeed");
            int32_t* op = add<int32_t>("veh_departspeed_proc", SB2_OFF(veh_departspeed_proc), nV, false, 0);
            double* ov = add<double>("veh_departspeed", SB2_OFF(veh_departspeed), nV, false, 0.);
            const int defProc = ff.scalar<int32_t>("opt_default_departspeed_proc", DSD_DEFAULT);
            const double defSpeed = ff.scalar<double>("opt_default_departspeed", 0.);
            for (int v = 0; v < nV; v++) {
                op[v] = sp[v];
                ov[v] = sv[v];
                if (op[v] == DSD_DEFAULT) {
                    op[v] = defProc;
                    ov[v] = defProc == DSD_GIVEN ? defSpeed : 0.;
                }
                if (op[v] == DSD_RANDOM || op[v] == DSD_GIVEN_VEHROUTE) {
                    throw std::runtime_error("sumob200: departSpeed random is not supported");
                }
                const int pp = h.veh_departpos_proc[v];
                if (pp != DPD_DEFAULT && pp != DPD_GIVEN && pp != DPD_BASE) {
                    throw std::runtime_error("sumob200: departPos must be 'base' or a number");
                }
            }
            // depart lane (MSEdge::getDepartLane, MSEdge.cpp:677-751): first lane / given index
            const int32_t* lp = ff.ptr<int32_t>("veh_departlane_proc");
            const int32_t* lv = ff.ptr<int32_t>("veh_departlane");
            int32_t* dl = add<int32_t>("veh_departlane", SB2_OFF(veh_departlane), nV, false, 0);
            const int defLaneProc = ff.scalar<int32_t>("opt_default_departlane_proc", DLD_DEFAULT);
            const int defLane = ff.scalar<int32_t>("opt_default_departlane", 0);
            for (int v = 0; v < nV; v++) {
                int proc = lp[v], idx = lv[v];
                if (proc == DLD_DEFAULT) {
                    proc = defLaneProc;
                    idx = defLane;
                }
                const int r = h.veh_route[v];
                const int e0 = h.route_edges[h.route_begin[r]];
                const int nl = h.edge_lane_begin[e0 + 1] - h.edge_lane_begin[e0];
                if (proc == DLD_DEFAULT || proc == DLD_FIRST_ALLOWED) {
                    idx = 0;
                } else if (proc != DLD_GIVEN) {
                    // free / allowed / best: every choice is the only lane of a single-lane edge; "random" draws from the RNG
                    if (nl != 1 || proc == 2) {
                        throw std::runtime_error("sumob200: departLane must be 'first' or an index on multi-lane edges");
                    }
                    idx = 0;
                }
                if (idx >= nl) {
                    throw std::runtime_error("sumob200: departLane index out of range");
                }
                dl[v] = h.edge_lane_begin[e0] + idx;
                // a vehicle inserted with its back behind the lane's start laps into the predecessor lanes (MSVehicle::computeFurtherLanes
                // at insertion, MSVehicle.cpp:5823) and MSLane::getFollowersOnConsecutive then asks the junction behind it for link
                // leaders (MSLane.cpp:3951-3975): not modelled.  On a lane without a predecessor (network fringe) there is nothing
                // behind: the back position is simply negative
                if (h.veh_departpos_proc[v] == DPD_GIVEN) {
                    double pos = h.veh_departpos[v];
                    if (pos < 0.) {
                        pos += h.lane_length[dl[v]];
                    }
                    if (pos < h.vt_length[h.veh_type[v]] && h.lane_pred_logical[dl[v]] >= 0) {
                        throw std::runtime_error("sumob200: departPos closer to the lane's start than the vehicle is long is not supported (on a lane with a predecessor)");
                    }
                }
            }
            // departure queues per lane, in load order
            int32_t* qb = add<int32_t>("dq_begin", SB2_OFF(dq_begin), (size_t)nL + 1, false, 0);
            int32_t* qv = add<int32_t>("dq_veh", SB2_OFF(dq_veh), nV, false, 0);
            int32_t* qh = SB2_DYN(int32_t, dq_head, nL, 0);
            for (int v = 0; v < nV; v++) {
                qb[dl[v] + 1]++;
            }
            for (int l = 0; l < nL; l++) {
                qb[l + 1] += qb[l];
                qh[l] = qb[l];
            }
            std::vector<int32_t> fill(qb, qb + nL);
            for (int v = 0; v < nV; v++) {
                const int slot = fill[dl[v]]++;
                qv[slot] = v;
                // MSInsertionControl tries vehicles in departure order: within a departure lane the slots must be sorted
                if (slot > qb[dl[v]] && h.veh_depart[qv[slot - 1]] > h.veh_depart[v]) {
                    throw std::runtime_error("sumob200: vehicles must be sorted by departure time");
                }
            }
        }
        // ---------------- vehicle state
        SB2_DYN(int32_t, veh_lane, nV, -1); SB2_DYN(int32_t, veh_routepos, nV, 0);
        SB2_DYN(double, veh_pos, nV, 0.); SB2_DYN(double, veh_speed, nV, 0.); SB2_DYN(double, veh_prevspeed, nV, 0.);
        SB2_DYN(double, veh_accel, nV, 0.); SB2_DYN(double, veh_backpos, nV, 0.); SB2_DYN(double, veh_lastdist, nV, 0.);
        SB2_DYN(int64_t, veh_waiting, nV, (int64_t)0); SB2_DYN(int64_t, veh_since_startup, nV, (int64_t)86400000);
        SB2_DYN(int64_t, veh_jentry, nV, (int64_t)SB2_TIME_MAX); SB2_DYN(int64_t, veh_jentry_ny, nV, (int64_t)SB2_TIME_MAX);
        SB2_DYN(int64_t, veh_jconflict, nV, (int64_t)SB2_TIME_MAX);
        SB2_DYN(uint8_t, veh_havetowait, nV, (uint8_t)0); SB2_DYN(uint8_t, veh_moved, nV, (uint8_t)0);
        SB2_DYN(int64_t, veh_depart_time, nV, (int64_t)-1);
        SB2_DYN(double, veh_odometer, nV, 0.); SB2_DYN(double, veh_timeloss, nV, 0.); SB2_DYN(double, veh_stopspeed, nV, SB2_DBL_MAX);
        SB2_DYN(int32_t, veh_ahead, nV, -1); SB2_DYN(int32_t, veh_behind, nV, -1); SB2_DYN(int32_t, veh_next_incoming, nV, -1);
        SB2_DYN(int32_t, veh_nfurther, nV, 0); SB2_DYN(int32_t, veh_further, (size_t)nV * MAX_FURTHER, -1);
        SB2_DYN(double, veh_len_in_front, nV, 0.);
        SB2_DYN(int32_t, tp_list, nV, -1); SB2_DYN(SUMOTime, veh_tp_proceed, nV, (SUMOTime)-1);
        SB2_DYN(uint8_t, edge_delayed, h.nE, (uint8_t)0);
        {
            int32_t* ext = SB2_DYN(int32_t, veh_ext, nV, 0);
            int32_t* inv = SB2_DYN(int32_t, ext2slot, nV, 0);
            for (int v = 0; v < nV; v++) {
                ext[v] = v;
                inv[v] = v;
            }
        }
        SB2_DYN(int32_t, veh_qlane, nV, -1); SB2_DYN(uint8_t, veh_deferred, nV, (uint8_t)0); SB2_DYN(int32_t, deferred_list, nV, -1);
        SB2_DYN(int32_t, lane_arrived, nL, 0);
        SB2_DYN(SUMOTime, veh_trip_wait, nV, 0); SB2_DYN(int32_t, veh_trip_waitcount, nV, 0); SB2_DYN(uint8_t, veh_trip_amwaiting, nV, 0);
        SB2_DYN(double, veh_trip_routelen, nV, 0.); SB2_DYN(double, veh_trip_departpos, nV, 0.); SB2_DYN(double, veh_trip_departspeed, nV, 0.);
        SB2_DYN(SUMOTime, veh_arrival_time, nV, -1); SB2_DYN(int32_t, veh_arrival_lane, nV, -1);
        {
            // emission device: per-vehicle flag + 7 accumulators; a vehicle with the device needs a supported (HBEFA3/HBEFA4) model
            bool anyDevice = false;
            if (ff.has("veh_emissions")) {
                uint8_t* flag = SB2_COPY(uint8_t, veh_emissions);
                for (int v = 0; v < nV; v++) {
                    if (flag[v]) {
                        anyDevice = true;
                        anyEmissionDevice = true;
                        if (!haveEmissionModel || h.vt_em_model[h.veh_type[v]] == 0) {
                            throw std::runtime_error("sumob200: emission device on a vehicle whose emission class is not HBEFA3/HBEFA4 (gasoline/diesel)");
                        }
                    }
                }
            } else {
                SB2_DYN(uint8_t, veh_emissions, nV, 0);
            }
            SB2_DYN(double, veh_em, anyDevice ? (size_t)nV * 7 : 1, 0.);
        }
        // drive items per vehicle: 8 covers the look-ahead on ordinary networks (and keeps the 1 M-vehicle bench at 2.2 GB); a network
        // of very short lanes (more than 8 links inside the braking distance) raises SUMOB200_F_DRIVE_ITEM_OVERFLOW -- the scenario
        // can then ask for more with the scalar drive_item_capacity (<= 64)
        h.max_items = std::min(64, std::max(2, ff.scalar<int32_t>("drive_item_capacity", MAX_DRIVE_ITEMS)));
        const size_t nDI = (size_t)nV * h.max_items;
        for (int b = 0; b < 2; b++) {
            SB2_DYN2(int32_t, veh_ndi, b, nV, 0);
            SB2_DYN2(int32_t, veh_order, b, nV, -1);
            SB2_DYN2(int32_t, di_link, b, nDI, -1); SB2_DYN2(uint8_t, di_request, b, nDI, (uint8_t)0); SB2_DYN2(uint8_t, di_hadstopped, b, nDI, (uint8_t)0);
            SB2_DYN2(double, di_vpass, b, nDI, 0.); SB2_DYN2(double, di_vwait, b, nDI, 0.); SB2_DYN2(double, di_arrspeed, b, nDI, 0.);
            SB2_DYN2(double, di_arrspeedbraking, b, nDI, 0.); SB2_DYN2(double, di_dist, b, nDI, 0.); SB2_DYN2(double, di_accelv, b, nDI, 0.);
            SB2_DYN2(double, di_avail, b, nDI, 0.);
            SB2_DYN2(int64_t, di_arrtime, b, nDI, (int64_t)0);
            SB2_DYN2(double, di_regspeed, b, nDI, 0.); SB2_DYN2(int64_t, di_regwaiting, b, nDI, (int64_t)0);
            {
                AppRec none;
                none.next = -1; none.veh = -1; none.arrtime = 0; none.leavetime = 0; none.arrspeedbraking = 0.;
                SB2_DYN2(AppRec, ap_rec, b, nDI, none);
            }
        }
        h.cur = 0;
        {
            double* f = SB2_DYN(double, fleet, 3, 0.);
            double* g = SB2_DYN(double, fleet_next, 3, 0.);
            f[0] = g[0] = ff.scalar<double>("opt_max_speed_factor", 1.);
            f[1] = g[1] = ff.scalar<double>("opt_min_deceleration", 4.5);
            f[2] = g[2] = ff.scalar<double>("opt_max_mingap", 0.);
        }
        // ---------------- E1 induction loops (optional arrays e1_lane, e1_pos, e1_length; sb2_e1.h)
        {
            const int nD = ff.has("e1_lane") ? (int)ff.count("e1_lane") : 0;
            if (nD > 0 && (!ff.has("e1_pos") || (int)ff.count("e1_pos") != nD || (ff.has("e1_length") && (int)ff.count("e1_length") != nD))) {
                throw std::invalid_argument("sumob200: e1_lane / e1_pos / e1_length must have the same length");
            }
            h.nE1 = nD;
            int32_t* lb = add<int32_t>("lane_e1_begin", SB2_OFF(lane_e1_begin), (size_t)nL + 1, false, 0);
            int32_t* of = add<int32_t>("e1_of_lane", SB2_OFF(e1_of_lane), (size_t)nD, false, 0);
            int32_t* el = add<int32_t>("e1_lane", SB2_OFF(e1_lane), (size_t)(nD > 0 ? nD : 1), false, 0);
            double* dp = add<double>("e1_pos", SB2_OFF(e1_pos), (size_t)nD, false, 0.);
            double* de = add<double>("e1_end", SB2_OFF(e1_end), (size_t)nD, false, 0.);
            if (nD > 0) {
                checkIndex(ff, "e1_lane", 0, nL);
                const int32_t* dl = ff.ptr<int32_t>("e1_lane");
                const double* pp = ff.ptr<double>("e1_pos");
                const double* pl = ff.has("e1_length") ? ff.ptr<double>("e1_length") : nullptr;
                for (int d = 0; d < nD; d++) {
                    // NLDetectorBuilder::getPositionChecking: a negative position counts from the lane's end
                    double pos = pp[d];
                    if (pos < 0) {
                        pos += h.lane_length[dl[d]];
                    }
                    const double len = pl != nullptr ? pl[d] : 0.;
                    if (pos < 0 || len < 0 || pos + len > h.lane_length[dl[d]]) {
                        throw std::invalid_argument("sumob200: an E1 detector does not lie on its lane");
                    }
                    el[d] = dl[d];
                    dp[d] = pos;
                    de[d] = pos + len;
                    lb[dl[d] + 1]++;
                }
                for (int l = 0; l < nL; l++) {
                    lb[l + 1] += lb[l];
                }
                std::vector<int32_t> fill(lb, lb + nL);
                for (int d = 0; d < nD; d++) {
                    of[fill[dl[d]]++] = d;
                }
            }
            const size_t nR = nD > 0 ? (size_t)nV * MAX_E1_REMINDERS : 1;
            SB2_DYN(int32_t, veh_e1_det, nR, -1); SB2_DYN(double, veh_e1_off, nR, 0.); SB2_DYN(double, veh_e1_entry, nR, E1_NOT_ON);
            SB2_DYN(int32_t, e1_entered, (size_t)nD, 0);
            SB2_DYN(int64_t, veh_sim_seen, (nD > 0 || ff.has("e2_lane")) ? (size_t)nV : 1, (int64_t)-1);
            // room for the passages of one output interval: every vehicle over a few detectors, or every detector a few thousand times
            h.e1_event_cap = nD > 0 ? (int)std::min<size_t>((size_t)1 << 26, std::max<size_t>((size_t)4 * nV + 65536, (size_t)64 * nD))
                                    : 1;
            if (ff.has("e1_event_capacity")) {
                h.e1_event_cap = std::max(1, ff.scalar<int32_t>("e1_event_capacity", 1));
            }
            E1Event none;
            none.det = -1; none.veh = -1; none.entry = 0; none.leave = 0; none.length = 0; none.leftEarly = 0; none.pad = 0;
            SB2_DYN(E1Event, e1_events, (size_t)h.e1_event_cap, none);
        }
        // ---------------- E2 lane area detectors on one lane (optional arrays e2_lane, e2_start, e2_length, e2_jam_*; sb2_e2.h)
        {
            const int nD = ff.has("e2_lane") ? (int)ff.count("e2_lane") : 0;
            for (const char* name : {"e2_start", "e2_length", "e2_jam_time", "e2_jam_speed", "e2_jam_dist"}) {
                if (nD > 0 && (!ff.has(name) || (int)ff.count(name) != nD)) {
                    throw std::invalid_argument(std::string("sumob200: ") + name + " must be given for every E2 detector");
                }
            }
            h.nE2 = nD;
            const size_t n1 = nD > 0 ? (size_t)nD : 1;
            int32_t* lb = add<int32_t>("lane_e2_begin", SB2_OFF(lane_e2_begin), (size_t)nL + 1, false, 0);
            int32_t* of = add<int32_t>("e2_of_lane", SB2_OFF(e2_of_lane), n1, false, 0);
            int32_t* la = add<int32_t>("e2_lane", SB2_OFF(e2_lane), n1, false, 0);
            double* ds = add<double>("e2_start", SB2_OFF(e2_start), n1, false, 0.);
            double* dl = add<double>("e2_length", SB2_OFF(e2_length), n1, false, 1.);
            double* js = add<double>("e2_jam_speed", SB2_OFF(e2_jam_speed), n1, false, 0.);
            double* jd = add<double>("e2_jam_dist", SB2_OFF(e2_jam_dist), n1, false, 0.);
            int64_t* jt = add<int64_t>("e2_jam_time", SB2_OFF(e2_jam_time), n1, false, (int64_t)0);
            if (nD > 0) {
                checkIndex(ff, "e2_lane", 0, nL);
                const int32_t* fl = ff.ptr<int32_t>("e2_lane");
                for (int d = 0; d < nD; d++) {
                    la[d] = fl[d];
                    ds[d] = ff.ptr<double>("e2_start")[d];
                    dl[d] = ff.ptr<double>("e2_length")[d];
                    js[d] = ff.ptr<double>("e2_jam_speed")[d];
                    jd[d] = ff.ptr<double>("e2_jam_dist")[d];
                    jt[d] = ff.ptr<int64_t>("e2_jam_time")[d];
                    if (ds[d] < 0 || dl[d] <= 0 || ds[d] + dl[d] > h.lane_length[la[d]] + SB2_NUMERICAL_EPS) {
                        throw std::invalid_argument("sumob200: an E2 detector does not lie on its lane");
                    }
                    lb[la[d] + 1]++;
                }
                for (int l = 0; l < nL; l++) {
                    lb[l + 1] += lb[l];
                }
                std::vector<int32_t> fill(lb, lb + nL);
                for (int d = 0; d < nD; d++) {
                    of[fill[la[d]]++] = d;
                }
            }
            const size_t nR = nD > 0 ? (size_t)nV * MAX_E2_REMINDERS : 1;
            SB2_DYN(int32_t, veh_e2_det, nR, -1); SB2_DYN(double, veh_e2_off, nR, 0.); SB2_DYN(int32_t, veh_e2_entered, nR, 0);
            // one note per vehicle, detector and step; a vehicle is rarely on more than two detectors at once (overflow raises the flag)
            h.e2_note_cap = nD > 0 ? (int)std::min<size_t>((size_t)1 << 26, std::min((size_t)nV * MAX_E2_REMINDERS, (size_t)2 * nV + 65536)) : 1;
            if (ff.has("e2_note_capacity")) {
                h.e2_note_cap = std::max(1, ff.scalar<int32_t>("e2_note_capacity", 1));
            }
            E2Note none;
            none.next = -1; none.veh = -1; none.dist = 0; none.timeOn = 0; none.lenOn = 0; none.timeLoss = 0; none.speed = 0; none.halt = 0; none.ihalt = 0;
            for (int b = 0; b < 2; b++) {
                SB2_DYN2(E2Note, e2_notes, b, (size_t)h.e2_note_cap, none);
                SB2_DYN2(int32_t, e2_head, b, n1, -1);
            }
            SB2_DYN(double, e2_acc_d, n1 * E2D_N, 0.);
            SB2_DYN(int64_t, e2_acc_i, n1 * E2I_N, (int64_t)0);
            SB2_DYN(int32_t, e2_cnt, n1 * E2C_N, 0);
        }
        // ---------------- edgeData / laneData accumulators (optional scalar md_enabled; sb2_meandata.h)
        {
            h.md_on = ff.scalar<int32_t>("md_enabled", 0) != 0 ? 1 : 0;
            h.md_halt_speed = ff.scalar<double>("md_halt_speed", SB2_POSITION_EPS);
            const size_t n = h.md_on ? (size_t)nL : 1;
            SB2_DYN(double, md_sums, n * MD_NSUMS, 0.);
            SB2_DYN(double, md_minlen, n, SB2_DBL_MAX);
            SB2_DYN(int32_t, md_counts, n * MD_NCOUNTS, 0);
        }
        SB2_DYN(int32_t, err_flags, 1, 0);
        SB2_DYN(int32_t, counters, 16, 0);
    }
};

}  // namespace sb2
