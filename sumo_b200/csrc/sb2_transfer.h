/* sb2_transfer.h -- teleporting vehicles: MSVehicleTransfer::checkInsertions and MSLane::freeInsertion.
 *
 * Restates src/microsim/MSVehicleTransfer.cpp:97-217 (checkInsertions; no parking, no jumps), src/microsim/MSLane.cpp:516-602
 * (freeInsertion with NOTIFICATION_TELEPORT), src/microsim/MSEdge.cpp:547-575 (getFreeLane), 991-1028 (getMeanSpeed), 1065-1071
 * (getCurrentTravelTime), 122-127 (myEmptyTraveltime), src/microsim/MSLane.cpp:3474-3492 (getMeanSpeed), 2929-2941
 * (getMissingRearGap).  How a vehicle gets here: sb2_lane.h veh_transfer_add (waited longer than time-to-teleport, or collided).
 *
 * The reference handles the waiting vehicles one after the other in numerical-id order and every success changes what the
 * next one sees, so this runs as ONE thread (k_transfer); the list is empty in almost every step.
 */
#pragma once
#include "sb2_insert.h"

namespace sb2 {

// MSLane::getMeanSpeed (no stops on this path)
SB2_HD double lane_mean_speed(const Ctx& c, int l) {
    if (c.lane_nveh[l] == 0) {
        return c.lane_speed[l];
    }
    double v = 0;
    int num = 0;
    for (int x = c.lane_last[l]; x >= 0; x = c.veh_ahead[x]) {
        v += c.veh_speed[x];
        num++;
    }
    return v / num;
}

// what libsumo::Lane::getLastStepVehicleNumber / HaltingNumber / MeanSpeed / Occupancy / getWaitingTime return for lane l
// (src/libsumo/Lane.cpp:209-272, MSLane::getMeanSpeed / getNettoOccupancy / getFractionalVehicleLength, MSLane.cpp:3415-3494).
// The reference keeps the netto length sum incrementally; here the lengths of the queue are added rear to front.
struct LaneStats {   // == sumob200_lane_stats
    int32_t vehicleNumber, haltingNumber;
    double meanSpeed, occupancy, waitingTime;
};
SB2_HD LaneStats lane_stats(const Ctx& c, int l) {
    LaneStats r;
    r.vehicleNumber = 0; r.haltingNumber = 0; r.waitingTime = 0.;
    double netto = 0.;
    for (int x = c.lane_last[l]; x >= 0; x = c.veh_ahead[x]) {
        r.vehicleNumber++;
        if (c.veh_speed[x] < SB2_HALTING_SPEED) {
            r.haltingNumber++;
        }
        netto += veh_length(c, x);
        r.waitingTime += STEPS2TIME(c.veh_waiting[x]);
    }
    r.meanSpeed = lane_mean_speed(c, l);
    double fractions = 0;
    for (int i = 0; i < c.lane_npartial[l]; i++) {
        const int p = c.lane_partial[(size_t)l * MAX_PARTIAL + i];
        fractions += c.lane_length[l] - veh_back_pos_on_lane(c, p, l);   // getFractionalVehicleLength(false)
    }
    const int last = c.lane_last[l];
    if (last >= 0 && c.veh_pos[last] < veh_length(c, last)) {
        fractions -= veh_length(c, last) - c.veh_pos[last];
    }
    r.occupancy = (netto + fractions) / c.lane_length[l];
    return r;
}

// MSEdge::getCurrentTravelTime(minSpeed)
SB2_HD double edge_current_travel_time(const Ctx& c, int e, double minSpeed) {
    const int l0 = c.edge_lane_begin[e];
    const double length = c.lane_length[l0];
    if (!c.edge_delayed[e]) {
        return length / MAX2(c.lane_speed[l0], SB2_NUMERICAL_EPS);   // myEmptyTraveltime (no TLS / minor penalties in the microsim default)
    }
    double v = 0;
    double total = 0;
    for (int l = l0; l < c.edge_lane_begin[e + 1]; l++) {
        int n = c.lane_nveh[l];
        if (n == 0) {
            n = 1;
        }
        v += n * lane_mean_speed(c, l);
        total += n;
    }
    return length / MAX2(minSpeed, v / total);
}

// MSEdge::getFreeLane(allowed, vclass, departPos = 0): all lanes of the edge that continue to nextEdge (or all lanes); the least
// occupied one, overruled by the one with the largest gap behind its last vehicle
SB2_HD int edge_free_lane(const Ctx& c, int e, int nextEdge) {
    int res = -1, resByGap = -1;
    double largestGap = 0;
    double leastOccupancy = SB2_DBL_MAX;
    for (int l = c.edge_lane_begin[e]; l < c.edge_lane_begin[e + 1]; l++) {
        if (nextEdge >= 0 && lane_succ_link(c, l, nextEdge) < 0) {
            continue;
        }
        // MSLane::getBruttoOccupancy (MSLane.cpp:3434-3445) with getFractionalVehicleLength (MSLane.cpp:3415-3431)
        double fractions = 0;
        for (int i = 0; i < c.lane_npartial[l]; i++) {
            fractions += c.lane_length[l] - veh_back_pos_on_lane(c, c.lane_partial[(size_t)l * MAX_PARTIAL + i], l);
        }
        const int last = c.lane_last[l];
        if (last >= 0 && c.veh_pos[last] < veh_length(c, last)) {
            fractions -= veh_length(c, last) - c.veh_pos[last];
        }
        const double occ = c.lane_brutto[l] + fractions;
        const double occupancy = MIN2(1., occ / c.lane_length[l]);
        if (occupancy < leastOccupancy) {
            res = l;
            leastOccupancy = occupancy;
        }
        const double lastGap = (last >= 0 ? c.veh_pos[last] : c.lane_length[l]) - 0.;
        if (lastGap > largestGap) {
            largestGap = lastGap;
            resByGap = l;
        }
    }
    return resByGap >= 0 ? resByGap : res;
}

// MSLane::freeInsertion(veh, mspeed, 0, NOTIFICATION_TELEPORT)
SB2_HD bool lane_freeInsertion(const Ctx& c, int l, int v, double mspeed) {
    CF cf = make_cf(c, c.veh_type[v]);
    cf.desSpeed = lane_vehicle_max_speed(c, l, v);   // setTentativeLaneAndPosition(this, ...): veh->getLane() is this lane (IDM reads it)
    const double maxPos = c.lane_length[l];
    const double minPos = MIN2(maxPos, cf.length());
    const double egoSpeed = c.veh_speed[v];
    if (c.lane_nveh[l] == 0) {
        // ensure sufficient gap to followers on predecessor lanes: getMissingRearGap(&veh, backOffset, mspeed)
        const double backOffset = minPos - cf.length();
        const FollowerInfo fi = lane_followers_on_consecutive(c, l, v, cf.length(), cf.decel(), backOffset, egoSpeed, false);
        double missingRearGap = 0;
        if (fi.veh >= 0) {
            const CF fcf = make_cf(c, c.veh_type[fi.veh]);
            missingRearGap = cf_getSecureGap(fcf, c.veh_speed[fi.veh], mspeed, cf.decel()) - fi.dist;
        }
        if (missingRearGap > 0) {
            if (minPos + missingRearGap <= maxPos) {
                return lane_isInsertionSuccess(c, l, v, mspeed, minPos + missingRearGap, true, true, true);
            }
            return false;
        }
        return lane_isInsertionSuccess(c, l, v, mspeed, minPos, true, true, true);
    } else {
        // behind the last vehicle ("reproduction of bogus old behavior", MSLane.cpp:547: `leader` is myVehicles.back(), the FRONT-most one)
        const int leader = c.lane_first[l];
        const double leaderPos = veh_back_pos_on_lane(c, leader, l);
        const double speed = c.veh_speed[leader];
        const double frontGapNeeded = cf_getSecureGap(cf, speed, c.veh_speed[leader], c.vt_decel[c.veh_type[leader]]) + cf.minGap();
        if (leaderPos >= frontGapNeeded) {
            const double tspeed = MIN2(cf_insertionFollowSpeed(cf, mspeed, frontGapNeeded, c.veh_speed[leader], c.vt_decel[c.veh_type[leader]]), mspeed);
            if (lane_isInsertionSuccess(c, l, v, tspeed, minPos, true, true, true)) {
                return true;
            }
        }
    }
    // go through the lane, look for free positions (starting after the last vehicle)
    for (int follower = c.lane_last[l]; follower >= 0; follower = c.veh_ahead[follower]) {
        int leader = c.veh_ahead[follower];
        if (leader < 0 && c.lane_npartial[l] > 0) {
            leader = c.lane_partial[(size_t)l * MAX_PARTIAL];
        }
        double speed = mspeed;
        if (leader >= 0) {
            speed = MIN2(c.veh_speed[leader], mspeed);
        }
        double frontMax = maxPos;
        if (leader >= 0) {
            const double leaderRearPos = veh_back_pos_on_lane(c, leader, l);
            const double frontGapNeeded = cf_getSecureGap(cf, speed, c.veh_speed[leader], c.vt_decel[c.veh_type[leader]]) + cf.minGap();
            frontMax = MIN2(maxPos, leaderRearPos - frontGapNeeded);
        }
        const CF fcf = make_cf(c, c.veh_type[follower]);
        const double followPos = c.veh_pos[follower] + fcf.minGap();
        const double backGapNeeded = cf_getSecureGap(fcf, c.veh_speed[follower], egoSpeed, cf.decel());
        const double backMin = followPos + backGapNeeded + cf.length();
        if (frontMax > minPos && backMin + SB2_POSITION_EPS < frontMax) {
            if (lane_isInsertionSuccess(c, l, v, speed, backMin + SB2_POSITION_EPS, true, true, true)) {
                return true;
            }
        }
    }
    return false;
}

// MSVehicleTransfer::checkInsertions(time)
SB2_HDN void transfer_checkInsertions(const Ctx& c) {
    const int n = c.counters[10];
    int32_t* list = c.tp_list;
    // std::sort(vehInfos) by numerical id; the list is almost sorted (only this step's additions are out of place)
    for (int i = 1; i < n; i++) {
        const int x = list[i];
        int j = i - 1;
        while (j >= 0 && c.veh_numid[list[j]] > c.veh_numid[x]) {
            list[j + 1] = list[j];
            j--;
        }
        list[j + 1] = x;
    }
    const SUMOTime time = c.t;
    int keep = 0;
    for (int i = 0; i < n; i++) {
        const int v = list[i];
        const int e = veh_succ_edge(c, v, 0);
        const int nextEdge = veh_succ_edge(c, v, 1);
        if (!lane_auth(c, c.edge_lane_begin[e])) {
            list[keep++] = v;   // moved on into another rank's part of the network: handed over with the next halo exchange
            continue;
        }
        const int l = edge_free_lane(c, e, nextEdge);
        const double mspeed = l >= 0 ? MIN2(c.lane_speed[l], veh_max_speed(c, v)) : 0.;
        if (l >= 0 && lane_freeInsertion(c, l, v, mspeed)) {
            continue;   // "ends teleporting"
        }
        // could not insert: maybe proceed in virtual space
        if (c.veh_tp_proceed[v] < 0) {
            c.veh_tp_proceed[v] = time + TIME2STEPS(edge_current_travel_time(c, e, 1.));   // TeleportMinSpeed = 1
        } else if (c.veh_tp_proceed[v] < time) {
            if (nextEdge < 0) {
                // "teleports beyond arrival edge": leaveLane(NOTIFICATION_TELEPORT_ARRIVED), scheduleVehicleRemoval
                // (MSVehicleControl::removePending runs in the NEXT step's executeMovements: until then the vehicle still counts as
                // running in the summary output -- veh_moved 4, see summary_vehicle)
                veh_e1_leave_net(c, v);
                veh_e2_leave_net(c, v);
                md_count(c, c.edge_lane_begin[e], MD_LEFT);        // leaveLane(NOTIFICATION_TELEPORT_ARRIVED)
                md_count(c, c.edge_lane_begin[e], MD_TELEPORTED);
                c.veh_arrival_time[v] = time;
                c.veh_arrival_lane[v] = c.edge_lane_begin[e];
                c.veh_trip_amwaiting[v] = 3;   // NOTIFICATION_TELEPORT_ARRIVED for the trip record, removal pending until the next step
                c.veh_moved[v] = 4;
                c.counters[11] += 1;   // still counted as running in this step's result
                c.veh_lane[v] = -2;
                count_event(c, c.edge_lane_begin[e], 1);
                continue;
            }
            // leaveLane(NOTIFICATION_TELEPORT_CONTINUATION), enterLaneAtMove(first lane of the next edge, onTeleporting)
            const int cl = c.edge_lane_begin[e];
            c.veh_odometer[v] += c.lane_length[cl];
            c.veh_trip_routelen[v] += c.lane_length[cl];
            c.veh_routepos[v] += 1;
            c.veh_tp_proceed[v] = time + TIME2STEPS(edge_current_travel_time(c, e, 1.));
            veh_e1_leave_net(c, v);
            veh_e1_add_lane(c, v, c.edge_lane_begin[nextEdge], true, 0., -veh_length(c, v));
            veh_e2_leave_net(c, v);
            veh_e2_add_lane(c, v, c.edge_lane_begin[nextEdge], true, -veh_length(c, v));
            md_count(c, cl, MD_LEFT);                              // leaveLane(NOTIFICATION_TELEPORT_CONTINUATION)
            md_count(c, c.edge_lane_begin[nextEdge], MD_ENTERED);  // enterLaneAtMove(.., onTeleporting) -> notifyEnter(NOTIFICATION_TELEPORT)
        }
        list[keep++] = v;
    }
    c.counters[10] = keep;
}

}  // namespace sb2
