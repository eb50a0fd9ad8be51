/* sb2_veh.h -- per-vehicle accessors shared by the plan / execute / insertion code.
 * Reference: MSBaseVehicle.cpp:211-213 (getMaxSpeed), MSBaseVehicle.h:605 (getDesiredMaxSpeed), MSBaseVehicle.cpp:1234-1238
 * (getImpatience), MSLane.h:580-592 (getVehicleMaxSpeed), MSVehicle.cpp:5135-5239 (getBackPositionOnLane), :5248 (isFrontOnLane),
 * MSLane.cpp:2700-2715 (getLastAnyVehicle), :1477-1529 (getLastVehicleInformation with ego == nullptr, minPos == 0).
 */
#pragma once
#include "sb2_types.h"
#include "sb2_cf.h"

namespace sb2 {

SB2_HD double veh_length(const Ctx& c, int v) { return c.vt_length[c.veh_type[v]]; }
SB2_HD double veh_mingap(const Ctx& c, int v) { return c.vt_mingap[c.veh_type[v]]; }
SB2_HD double veh_length_with_gap(const Ctx& c, int v) { const int t = c.veh_type[v]; return c.vt_length[t] + c.vt_mingap[t]; }
SB2_HD double veh_max_speed(const Ctx& c, int v) {
    const int t = c.veh_type[v];
    return MIN2(c.vt_maxspeed[t], c.vt_desiredmaxspeed[t] * c.veh_speedfactor[v]);
}
SB2_HD double veh_desired_max_speed(const Ctx& c, int v) { return c.vt_desiredmaxspeed[c.veh_type[v]] * c.veh_speedfactor[v]; }
SB2_HD double lane_vehicle_max_speed(const Ctx& c, int l, int v, double vehMaxSpeed) {
    return MIN2(vehMaxSpeed, c.lane_speed[l] * c.veh_speedfactor[v]);
}
SB2_HD double lane_vehicle_max_speed(const Ctx& c, int l, int v) { return lane_vehicle_max_speed(c, l, v, veh_max_speed(c, v)); }

SB2_HD double veh_impatience(const Ctx& c, int v) {
    return MAX2(0., MIN2(1., c.vt_impatience[c.veh_type[v]]
                         + 0
                         + (c.timeToImpatience > 0 ? div_pos((double)c.veh_waiting[v], (double)c.timeToImpatience) : 0.)));
}

SB2_HD int veh_route_len(const Ctx& c, int v) { const int r = c.veh_route[v]; return c.route_begin[r + 1] - c.route_begin[r]; }
// MSBaseVehicle::succEdge(n): edge index or -1
SB2_HD int veh_succ_edge(const Ctx& c, int v, int n) {
    const int r = c.veh_route[v];
    const int i = c.veh_routepos[v] + n;
    return (i < c.route_begin[r + 1] - c.route_begin[r]) ? c.route_edges[c.route_begin[r] + i] : -1;
}

// MSVehicle::getBackPositionOnLane(lane) for the cases of the non-sublane, non-opposite model
SB2_HD double veh_back_pos_on_lane(const Ctx& c, int v, int lane) {
    (void)SB2_CHK(v, 0, c.nV, 6);
    const double len = veh_length(c, v);
    if (lane == c.veh_lane[v]) {
        return c.veh_pos[v] - len;
    }
    const int nf = c.veh_nfurther[v];
    const int32_t* f = c.veh_further + (size_t)v * MAX_FURTHER;
    if (nf > 0 && lane == f[nf - 1]) {
        return c.veh_backpos[v];
    }
    double leftLength = len - c.veh_pos[v];
    for (int i = 0; i < nf && leftLength > 0; i++) {
        leftLength -= c.lane_length[SB2_CHK(f[i], 0, c.nL, 7)];
        if (f[i] == lane) {
            return -leftLength;
        }
    }
    return c.veh_backpos[v];
}
// MSVehicle::getPositionOnLane(lane) = getBackPositionOnLane(lane, true) + length
SB2_HD double veh_pos_on_lane(const Ctx& c, int v, int lane) { return veh_back_pos_on_lane(c, v, lane) + veh_length(c, v); }
SB2_HD bool veh_front_on_lane(const Ctx& c, int v, int lane) { return c.veh_lane[v] == lane; }
// MSVehicle::getBackLane (MSVehicle.cpp:5045)
SB2_HD int veh_back_lane(const Ctx& c, int v) {
    const int nf = c.veh_nfurther[v];
    return nf > 0 ? c.veh_further[(size_t)v * MAX_FURTHER + nf - 1] : c.veh_lane[v];
}

// MSLane::getLastAnyVehicle
SB2_HD int lane_last_any(const Ctx& c, int l) {
    if (c.lane_last[l] >= 0) {
        return c.lane_last[l];
    }
    return c.lane_npartial[l] > 0 ? c.lane_partial[(size_t)l * MAX_PARTIAL] : -1;
}

// iterate vehicles and partial occupators of a lane in ascending position (MSLane::AnyVehicleIterator, downstream)
struct AnyVehIter {
    int veh;      // next entry of the lane queue (rear to front)
    int ip;       // next partial occupator
};
SB2_HD AnyVehIter any_begin(const Ctx& c, int l) { AnyVehIter it; it.veh = c.lane_last[l]; it.ip = 0; return it; }
SB2_HD int any_next(const Ctx& c, int l, AnyVehIter& it) {
    const int np = c.lane_npartial[l];
    const bool haveV = it.veh >= 0, haveP = it.ip < np;
    if (!haveV && !haveP) {
        return -1;
    }
    bool takeV;
    if (!haveP) {
        takeV = true;
    } else if (!haveV) {
        takeV = false;
    } else {
        const int p = c.lane_partial[(size_t)l * MAX_PARTIAL + it.ip];
        takeV = c.veh_pos[it.veh] < veh_pos_on_lane(c, p, l);   // MSLane.cpp:237
    }
    if (takeV) {
        const int r = it.veh;
        it.veh = c.veh_ahead[r];
        return r;
    }
    return c.lane_partial[(size_t)l * MAX_PARTIAL + it.ip++];
}

// link of lane `l` used to reach route edge `nextEdge` (MSLane::succLinkSec, MSLane.cpp:2733-2780; one lane per edge or
// first matching lane: bestLanes continuation == the only lane)
SB2_HD int lane_succ_link(const Ctx& c, int l, int nextEdge) {
    if (nextEdge < 0) {
        return -1;
    }
    if (lane_internal(c, l)) {
        return c.lane_link_begin[l];
    }
    for (int k = c.lane_link_begin[l]; k < c.lane_link_begin[l + 1]; k++) {
        const int to = c.link_to[k];
        if (to >= 0 && c.lane_edge[to] == nextEdge) {
            return k;
        }
    }
    return -1;
}

// executeMove appends the vehicle to the partial list of every lane its back still covers
SB2_HD void veh_register_partials(const Ctx& c, int v) {
    if (c.veh_lane[v] < 0 || c.veh_moved[v] == 2) {
        return;
    }
    const int nf = c.veh_nfurther[v];
    for (int i = 0; i < nf; i++) {
        const int l = c.veh_further[(size_t)v * MAX_FURTHER + i];
#if defined(__CUDA_ARCH__)
        const int slot = atomicAdd(&c.lane_npartial[l], 1);
#else
        const int slot = c.lane_npartial[l]++;
#endif
        if (slot < MAX_PARTIAL) {
            c.lane_partial[(size_t)l * MAX_PARTIAL + slot] = v;
        }
    }
}

}  // namespace sb2
