/* sumob200_flatfile.h -- the ".flat" container: a flattened SUMO scenario (network + vTypes + routes + departures)
 * as named plain arrays.  This is the data that crosses the drop-in boundary: whatever loads
 * the scenario (the reference's NLBuilder via integration/sumo_flatten.cpp, or this repo's grid tiler) fills these
 * arrays; the engine (libsumob200.so) only ever sees them.
 *
 * File layout (little endian):
 *   char     magic[8] = "SB2FLAT1"
 *   uint32   n_arrays
 *   n_arrays x { char name[40]; uint32 dtype; uint32 elem_size; uint64 count; uint8 data[count*elem_size]; pad to 8 }
 *
 * dtype: 0=u8 1=i32 2=i64 3=f64 4=u64
 *
 * Array names and meaning (all indices 0-based, -1 = none).  Reference accessors each array mirrors are cited.
 *  --- lanes (nL; ordered edge by edge, MSEdge::getAllEdges() order, lanes of one edge contiguous) ---
 *  lane_length f64       MSLane::getLength            lane_speed f64   MSLane::getSpeedLimit
 *  lane_width f64        MSLane::getWidth             lane_edge i32    index into edges
 *  lane_flags i32        bit0 internal edge, bit1 roundabout, bit2 fringe, bit3 edge has lane changer
 *  lane_numid i32        MSLane::getNumericalID (decides RNG bucket, MSLane.cpp:296)
 *  lane_link_begin i32[nL+1]   CSR into links (MSLane::getLinkCont order)
 *  lane_pred_logical i32 MSLane::getLogicalPredecessorLane   lane_pred_normal i32  getNormalPredecessorLane
 *  lane_entry_link i32   MSLane::getEntryLink (internal lanes)  lane_inc_link i32  getIncomingLanes()[0].viaLink (internal lanes)
 *  lane_inc_begin i32[nL+1], inc_lane i32, inc_link i32, inc_length f64   MSLane::getIncomingLanes (all lanes)
 *  lane_to_junction i32, lane_from_junction i32   MSEdge::getToJunction / getFromJunction
 *  lane_shape_begin i32[nL+1], shape_xy f64[2*nP]   MSLane::getShape
 *  lane_geomfactor f64   MSLane::getLengthGeometryFactor
 *  lane_name_begin i32[nL+1], lane_names u8       MSLane::getID
 *  --- edges (nE) ---
 *  edge_lane_begin i32[nE+1], edge_name_begin i32[nE+1], edge_names u8, edge_flags i32 (as lane_flags)
 *  --- links (nK; grouped by lane_link_begin) ---
 *  link_from i32 getLaneBefore  link_to i32 getLane  link_via i32 getViaLane  link_state u8 getState  link_dir u8 (char of LinkDirection)
 *  link_length f64 getLength    link_foevis f64 getFoeVisibilityDistance   link_radius f64 (myRadius, MSLink.cpp:553)
 *  link_flags i32  bit0 keepClear bit1 hasFoes bit2 amCont bit3 isEntryLink bit4 isConflictEntryLink bit5 isExitLink
 *                  bit6 isExitLinkAfterInternalJunction bit7 isInternalJunctionLink bit8 isIndirect bit9 havePedestrianCrossingFoe
 *                  bit10 isTurnaround bit11 amContOff
 *  link_index i32 getIndex   link_tl i32 (index into tls or -1)   link_tlindex i32 getTLIndex
 *  link_internal_before i32 getInternalLaneBefore   link_junction i32   link_entry i32 getCorrespondingEntryLink
 *  link_exit i32 getCorrespondingExitLink
 *  link_foelink_begin i32[nK+1], foelink i32                      getFoeLinks
 *  link_foelane_begin i32[nK+1], foelane i32, conf_lbc f64, conf_size f64, conf_flag i32, conf_foeidx i32   getFoeLanes/getConflicts
 *  link_response u64[4*nK]    MSJunctionLogic::getResponseFor(getIndex()) as 256-bit mask (only for links from internal lanes' junction logic)
 *  --- junctions (nJ) ---  junction_name_begin, junction_names
 *  --- traffic lights (static programs) ---
 *  tl_phase_begin i32[nTL+1], phase_dur i64, phase_state_begin i32[nPh+1], phase_states u8, tl_cur_phase i32, tl_next_switch i64
 *  tl_link_begin i32[nTL+1] (per tl: one entry per (tlIndex,link) pair), tllink_index i32, tllink_link i32
 *  --- vTypes (nT) ---
 *  vt_length, vt_mingap, vt_maxspeed, vt_desiredmaxspeed, vt_accel, vt_decel, vt_emergencydecel, vt_apparentdecel, vt_tau,
 *  vt_sigma, vt_impatience, vt_width (f64)   vt_cfmodel i32 (SUMO_TAG id of the model)  vt_startupdelay i64
 *  vt_idm_delta, vt_idm_stepping f64   vt_emission_class i32   vt_name_begin, vt_names
 *  --- routes (nR) ---  route_begin i32[nR+1], route_edges i32
 *  --- vehicles (nV; in insertion order = stable order by depart time, MSVehicleContainer) ---
 *  veh_numid i32 getNumericalID  veh_type i32  veh_route i32  veh_depart i64  veh_speedfactor f64 getChosenSpeedFactor
 *  veh_arrivalpos f64 getArrivalPos   veh_departpos_proc i32, veh_departpos f64, veh_departspeed_proc i32, veh_departspeed f64
 *  veh_departlane_proc i32, veh_departlane i32   veh_name_begin, veh_names   veh_idrank i32 (rank of id in std::string order)
 *  --- scalars (count 1) ---
 *  opt_delta_t i64, opt_begin i64, opt_end i64, opt_seed i64, opt_time_to_teleport i64, opt_time_to_impatience i64,
 *  opt_default_departspeed_proc i32 ... (see integration/sumo_flatten.cpp)
 */
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace sb2 {

enum FlatDType : uint32_t { FD_U8 = 0, FD_I32 = 1, FD_I64 = 2, FD_F64 = 3, FD_U64 = 4 };

struct FlatArray {
    uint32_t dtype = 0;
    uint32_t elemSize = 1;
    uint64_t count = 0;
    std::vector<uint8_t> bytes;
    template<class T> const T* as() const { return reinterpret_cast<const T*>(bytes.data()); }
    template<class T> T* as() { return reinterpret_cast<T*>(bytes.data()); }
};

class FlatFile {
public:
    std::map<std::string, FlatArray> arrays;

    template<class T> static uint32_t dtypeOf();

    template<class T> void put(const std::string& name, const std::vector<T>& v) {
        FlatArray a;
        a.dtype = dtypeOf<T>();
        a.elemSize = sizeof(T);
        a.count = v.size();
        a.bytes.resize(v.size() * sizeof(T));
        if (!v.empty()) {
            memcpy(a.bytes.data(), v.data(), a.bytes.size());
        }
        arrays[name] = std::move(a);
    }
    template<class T> void putScalar(const std::string& name, T v) {
        put<T>(name, std::vector<T>(1, v));
    }
    bool has(const std::string& name) const { return arrays.count(name) != 0; }
    const FlatArray& get(const std::string& name) const {
        auto it = arrays.find(name);
        if (it == arrays.end()) {
            throw std::runtime_error("flat file: missing array '" + name + "'");
        }
        return it->second;
    }
    template<class T> const T* ptr(const std::string& name) const {
        const FlatArray& a = get(name);
        if (a.elemSize != sizeof(T)) {
            throw std::runtime_error("flat file: array '" + name + "' has wrong element size");
        }
        return a.as<T>();
    }
    uint64_t count(const std::string& name) const { return get(name).count; }
    template<class T> T scalar(const std::string& name, T dflt) const {
        auto it = arrays.find(name);
        if (it == arrays.end() || it->second.count == 0) {
            return dflt;
        }
        return *it->second.as<T>();
    }

    void save(const std::string& path) const {
        FILE* f = fopen(path.c_str(), "wb");
        if (f == nullptr) {
            throw std::runtime_error("flat file: cannot write '" + path + "'");
        }
        fwrite("SB2FLAT1", 1, 8, f);
        uint32_t n = (uint32_t)arrays.size();
        fwrite(&n, 4, 1, f);
        for (const auto& kv : arrays) {
            char name[40];
            memset(name, 0, sizeof(name));
            strncpy(name, kv.first.c_str(), 39);
            fwrite(name, 1, 40, f);
            fwrite(&kv.second.dtype, 4, 1, f);
            fwrite(&kv.second.elemSize, 4, 1, f);
            fwrite(&kv.second.count, 8, 1, f);
            fwrite(kv.second.bytes.data(), 1, kv.second.bytes.size(), f);
            const size_t pad = (8 - kv.second.bytes.size() % 8) % 8;
            const char zeros[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            fwrite(zeros, 1, pad, f);
        }
        fclose(f);
    }

    void load(const std::string& path) {
        FILE* f = fopen(path.c_str(), "rb");
        if (f == nullptr) {
            throw std::runtime_error("flat file: cannot read '" + path + "'");
        }
        char magic[8];
        uint32_t n = 0;
        if (fread(magic, 1, 8, f) != 8 || memcmp(magic, "SB2FLAT1", 8) != 0 || fread(&n, 4, 1, f) != 1) {
            fclose(f);
            throw std::runtime_error("flat file: bad header in '" + path + "'");
        }
        for (uint32_t i = 0; i < n; i++) {
            char name[41];
            name[40] = 0;
            FlatArray a;
            if (fread(name, 1, 40, f) != 40 || fread(&a.dtype, 4, 1, f) != 1 || fread(&a.elemSize, 4, 1, f) != 1 || fread(&a.count, 8, 1, f) != 1) {
                fclose(f);
                throw std::runtime_error("flat file: truncated '" + path + "'");
            }
            a.bytes.resize(a.count * a.elemSize);
            if (!a.bytes.empty() && fread(a.bytes.data(), 1, a.bytes.size(), f) != a.bytes.size()) {
                fclose(f);
                throw std::runtime_error("flat file: truncated data in '" + path + "'");
            }
            const size_t pad = (8 - a.bytes.size() % 8) % 8;
            char skip[8];
            if (pad > 0 && fread(skip, 1, pad, f) != pad) {
                fclose(f);
                throw std::runtime_error("flat file: truncated pad in '" + path + "'");
            }
            arrays[name] = std::move(a);
        }
        fclose(f);
    }
};

template<> inline uint32_t FlatFile::dtypeOf<uint8_t>() { return FD_U8; }
template<> inline uint32_t FlatFile::dtypeOf<int32_t>() { return FD_I32; }
template<> inline uint32_t FlatFile::dtypeOf<int64_t>() { return FD_I64; }
template<> inline uint32_t FlatFile::dtypeOf<double>() { return FD_F64; }
template<> inline uint32_t FlatFile::dtypeOf<uint64_t>() { return FD_U64; }

}  // namespace sb2
