/* sb2_insert.h -- vehicle departure (one GPU thread per departure lane; candidates of a lane are tried in load order and
 * the first failure blocks the lane for the step, like MSEdge::myFailedInsertionMemory).
 *
 * Restates src/microsim/MSInsertionControl.cpp:127-190 (emitVehicles/tryInsert), src/microsim/MSEdge.cpp:803-922
 * (insertVehicle), src/microsim/MSLane.cpp:606-661 (getDepartSpeed), 691-814 (insertVehicle: departPos given/base),
 * 818-847 (checkFailure), 851-1420 (isInsertionSuccess; non-rail, no stops, no pedestrians, no shadow lane),
 * 1435-1472 (safeInsertionSpeed), 2929-2953 (getMissingRearGap, getMaximumBrakeDist), 3840-4073
 * (getFollowersOnConsecutive, single sublane), 456-477 (incorporateVehicle), src/microsim/MSLeaderInfo.cpp:434-483
 * (MSCriticalFollowerDistanceInfo::addFollower), src/microsim/MSVehicle.cpp:5793-5858 (enterLaneAtInsertion),
 * src/microsim/MSBaseVehicle.cpp:1415-1423 (basePos).
 */
#pragma once
#include "sb2_plan.h"

namespace sb2 {


// Departures of one step run one thread per departure lane, concurrently.  A vehicle that another lane's thread inserts in this
// very step is INVISIBLE to this thread (snapshot semantics: deterministic, independent of thread timing): lane_tryInsert
// publishes the new vehicle's state before it links it into the lane queue (fence in between), and every read of ANOTHER lane's
// queue here skips entries with veh_depart_time == now.  (Teleporting vehicles re-enter afterwards, one by one, and see everything: snapshot == false.  The reference inserts sequentially in load order, so there the later
// vehicle sees the earlier one; that only matters when two vehicles depart in the same step on lanes closer to each other than
// the insertion look-ahead -- brake gap downstream, getMaximumBrakeDist upstream.)
// The queue heads and the fields of a possibly new vehicle are read past the L1 (ld_coherent): the SM may hold the cache line of
// veh_depart_time / veh_ahead / veh_behind from BEFORE the other thread's insertion (slots are sorted by departure lane, so
// neighbouring lanes' vehicles share lines) -- a stale -1 there made a new vehicle look like an old one whose state (veh_lane
// == -1) was then used: one wrong read in ~3e5 insertions on the 1000 x 1000 / 100 m grid, a fault where lane_edge[-1] is unmapped.
SB2_HD bool ins_is_new(const Ctx& c, int v) { return ld_coherent(&c.veh_depart_time[v]) == c.t; }
// getLastAnyVehicle of another lane, without this step's departures
SB2_HD int ins_lane_last_any(const Ctx& c, int l, bool snapshot) {
    int last = snapshot ? ld_coherent(&c.lane_last[l]) : c.lane_last[l];
    while (snapshot && last >= 0 && ins_is_new(c, last)) {
        last = ld_coherent(&c.veh_ahead[last]);
    }
    if (last >= 0) {
        return last;
    }
    return c.lane_npartial[l] > 0 ? c.lane_partial[(size_t)l * MAX_PARTIAL] : -1;
}
// front-most vehicle of another lane, without this step's departures
SB2_HD int ins_lane_first(const Ctx& c, int l, bool snapshot) {
    int first = snapshot ? ld_coherent(&c.lane_first[l]) : c.lane_first[l];
    while (snapshot && first >= 0 && ins_is_new(c, first)) {
        first = ld_coherent(&c.veh_behind[first]);
    }
    return first;
}

struct FollowerInfo {
    int veh;
    double dist, missingGap;
};

// MSCriticalFollowerDistanceInfo::addFollower, single sublane; egoSpeed = ego->getSpeed() (0 before departure)
SB2_HD void fi_addFollower(const Ctx& c, FollowerInfo& fi, int veh, double egoSpeed, double egoDecel, double gap) {
    if (veh < 0) {
        return;
    }
    const CF fcf = make_cf(c, c.veh_type[veh]);
    const double requiredGap = cf_getSecureGap(fcf, c.veh_speed[veh], egoSpeed, egoDecel);
    const double missingGap = requiredGap - gap;
    if ((missingGap > fi.missingGap
            || (missingGap > 0 && gap < fi.dist)
            || (gap < 0 && fi.dist > 0))
            && !(gap > 0 && fi.dist < 0)
            && !(fi.missingGap > 0 && fi.dist < gap)) {
        fi.veh = veh;
        fi.dist = gap;
        fi.missingGap = missingGap;
    }
}

// MSLane::getMaximumBrakeDist (MSLane.cpp:2945-2953)
SB2_HD double lane_max_brake_dist(const Ctx& c, int l) {
    const double maxSpeed = c.lane_speed[l] * c.fleet[0];
    return MIN2(maxSpeed * maxSpeed * 0.5 / c.fleet[1] + c.fleet[2], 1000.0);
}

// MSLane::getFollowersOnConsecutive(ego, backOffset, false) for a vehicle about to be inserted on lane l at pos
SB2_HD FollowerInfo lane_followers_on_consecutive(const Ctx& c, int l, int ego, double egoLength, double egoDecel, double backOffset, double egoSpeed, bool snapshot) {
    FollowerInfo fi;
    fi.veh = -1; fi.dist = SB2_DBL_MAX; fi.missingGap = -SB2_DBL_MAX;
    const double egoPos = backOffset + egoLength;
    AnyVehIter it = any_begin(c, l);
    for (int veh = any_next(c, l, it); veh >= 0; veh = any_next(c, l, it)) {
        const double vpos = veh_pos_on_lane(c, veh, l);
        if (veh != ego && vpos < egoPos) {
            const double dist = backOffset - vpos - veh_mingap(c, veh);
            fi_addFollower(c, fi, veh, egoSpeed, egoDecel, dist);
        }
    }
    if (fi.veh < 0) {
        double searchDist = lane_max_brake_dist(c, l) - backOffset;
        // breadth-first search upstream
        struct Inc { int lane; int link; double length; };
        Inc queueA[24], queueB[24];
        int visited[64];
        int nVisited = 0;
        const int e = c.lane_edge[l];
        for (int x = c.edge_lane_begin[e]; x < c.edge_lane_begin[e + 1] && nVisited < 64; x++) {
            visited[nVisited++] = x;
        }
        Inc* toExamine = queueA;
        Inc* newFound = queueB;
        int nEx = 0, nNew = 0;
        for (int i = c.lane_inc_begin[l]; i < c.lane_inc_begin[l + 1] && nEx < 24; i++) {
            toExamine[nEx].lane = c.inc_lane[i]; toExamine[nEx].link = c.inc_link[i]; toExamine[nEx].length = c.inc_length[i];
            nEx++;
        }
        while (nEx != 0) {
            for (int q = 0; q < nEx; q++) {
                const int next = SB2_CHK(toExamine[q].lane, 0, c.nL, 38);
                const double ilen = toExamine[q].length;
                searchDist = MIN2(searchDist, lane_max_brake_dist(c, next) - backOffset);
                // getFirstVehicleInformation: front-most of vehicles and partial occupators / front-most vehicle with front on lane
                const int np = c.lane_npartial[next];
                const int firstFront = SB2_CHK(ins_lane_first(c, next, snapshot), -1, c.nV, 39);
                int first = np > 0 ? c.lane_partial[(size_t)next * MAX_PARTIAL + np - 1] : firstFront;
                int v = SB2_CHK(first == ego ? firstFront : first, -1, c.nV, 40);
                double agap = 0;
                if (v >= 0 && v != ego) {
                    if (!veh_front_on_lane(c, v, next)) {
                        agap = ilen - c.lane_length[next] + backOffset;
                        agap -= veh_mingap(c, v);
                        const bool differentEdge = c.lane_edge[SB2_CHK(c.veh_lane[v], 0, c.nL, 20)] != c.lane_edge[l];
                        if (agap > 0 && differentEdge) {
                            v = firstFront;
                            if (v >= 0 && v != ego) {
                                agap = ilen - c.veh_pos[v] + backOffset - veh_mingap(c, v);
                            } else {
                                v = -1;
                            }
                        } else if (differentEdge && fi.veh == v) {
                            agap = 0;
                        }
                    } else {
                        agap = ilen - c.veh_pos[v] + backOffset - veh_mingap(c, v);
                    }
                    fi_addFollower(c, fi, v, egoSpeed, egoDecel, agap);
                }
                if (ilen < searchDist) {
                    for (int j = c.lane_inc_begin[next]; j < c.lane_inc_begin[next + 1]; j++) {
                        const int jl = c.inc_lane[j];
                        bool seenBefore = false;
                        for (int x = 0; x < nVisited; x++) {
                            if (visited[x] == jl) {
                                seenBefore = true;
                                break;
                            }
                        }
                        if (!seenBefore) {   // MinorLinkMode::FOLLOW_ALWAYS
                            if (nVisited < 64) {
                                visited[nVisited++] = jl;
                            }
                            if (nNew < 24) {
                                newFound[nNew].lane = jl; newFound[nNew].link = c.inc_link[j]; newFound[nNew].length = c.inc_length[j] + ilen;
                                nNew++;
                            } else {
                                raise(c, ERR_LINKLEADER_OVERFLOW);
                            }
                        }
                    }
                }
            }
            Inc* tmp = toExamine; toExamine = newFound; newFound = tmp;
            nEx = nNew;
            nNew = 0;
        }
    }
    return fi;
}

// MSLane::checkFailure (MSLane.cpp:818-847); all InsertionCheck bits set, gEmergencyInsert false
SB2_HD bool ins_checkFailure(const CF& cf, double& speed, double& dist, double nspeed, bool patchSpeed) {
    if (nspeed < speed) {
        if (patchSpeed) {
            speed = MIN2(nspeed, speed);
            dist = cf_brakeGap(cf, speed) + cf.minGap();
        } else if (speed > 0) {
            return true;
        }
    }
    return false;
}

// MSLane::safeInsertionSpeed (MSLane.cpp:1435-1472) for one leader
SB2_HD double lane_safeInsertionSpeed(const Ctx& c, int l, const CF& cf, double seen, int leader, double speed) {
    double nspeed = speed;
    if (leader >= 0) {
        const double gap = veh_back_pos_on_lane(c, leader, l) + seen - cf.minGap();
        if (gap < 0) {
            return SB2_INVALID_SPEED;
        }
        nspeed = MIN2(nspeed, cf_insertionFollowSpeed(cf, speed, gap, c.veh_speed[leader], c.vt_decel[c.veh_type[leader]]));
    }
    return nspeed;
}

// MSLane::isInsertionSuccess + incorporateVehicle (MSLane.cpp:851-1420, 456-477); returns true if the vehicle was inserted.
// teleport = the vehicle re-enters from MSVehicleTransfer (notification NOTIFICATION_TELEPORT): it has departed before, keeps its
// route position, speed (ego->getSpeed() in the follower / link-leader checks) and trip records
SB2_HD bool lane_isInsertionSuccess(const Ctx& c, int l, int v, double speed, double pos, bool patchSpeed, bool patchSpeedSpecial, bool teleport) {
    CF cf = make_cf(c, c.veh_type[v]);
    cf.desSpeed = lane_vehicle_max_speed(c, l, v);   // setTentativeLaneAndPosition (MSLane.cpp:851): veh->getLane() == this lane
    const double egoSpeed = teleport ? c.veh_speed[v] : 0.;   // getSpeed() of a not yet departed vehicle is 0
    const int routePos0 = teleport ? c.veh_routepos[v] : 0;
    if (pos < 0 || pos > c.lane_length[l]) {
        pos = c.lane_length[l];
    }
    double seen = c.lane_length[l] - pos;
    double dist = cf_brakeGap(cf, speed) + cf.minGap();
    // leader on the insertion lane: getLastVehicleInformation(aVehicle, 0, pos)
    int leader = -1;
    {
        AnyVehIter it = any_begin(c, l);
        for (int veh = any_next(c, l, it); veh >= 0; veh = any_next(c, l, it)) {
            if (veh != v && MAX2(0.0, veh_pos_on_lane(c, veh, l)) >= pos) {
                leader = veh;
                break;
            }
        }
    }
    leader = SB2_CHK(leader, -1, c.nV, 32);
    const double nspeed = lane_safeInsertionSpeed(c, l, cf, -pos, leader, speed);
    if (nspeed == SB2_INVALID_SPEED || ins_checkFailure(cf, speed, dist, nspeed, patchSpeed)) {
        return false;
    }
    // continuation lanes
    Ego e;
    e.v = v; e.vt = c.veh_type[v]; e.lane = l; e.pos = pos; e.speed = egoSpeed;
    e.cf = cf; e.length = cf.length(); e.minGap = cf.minGap(); e.waiting = 0;
    int nRouteSuccs = 1;
    int currentLane = l;
    const int r = c.veh_route[v];
    const int routeLen = c.route_begin[r + 1] - c.route_begin[r];
    SUMOTime arrivalTime = c.t + TIME2STEPS(seen / MAX2(speed, SB2_HALTING_SPEED));
    int guard = 0;
    while (seen < dist && guard++ < 32) {
        const int nextEdge = routePos0 + nRouteSuccs < routeLen ? c.route_edges[c.route_begin[r] + routePos0 + nRouteSuccs] : -1;
        const int link = SB2_CHK(lane_succ_link(c, SB2_CHK(currentLane, 0, c.nL, 33), SB2_CHK(nextEdge, -1, c.nE, 34)), -1, c.nK, 35);
        if (link < 0) {
            if (c.lane_edge[currentLane] == c.route_edges[c.route_begin[r + 1] - 1]) {
                // reached the end of the route (mayContinue: no further route)
                if (c.veh_arrivalspeed[v] >= 0.) {
                    const double remaining = seen + c.veh_arrivalpos[v] - c.lane_length[currentLane];
                    const double fspeed = cf_freeSpeed(cf, speed, remaining, c.veh_arrivalspeed[v], true);
                    if (ins_checkFailure(cf, speed, dist, fspeed, patchSpeedSpecial)) {
                        return false;   // "arrival speed too low"
                    }
                }
            } else {
                if (ins_checkFailure(cf, speed, dist, cf_insertionStopSpeed(cf, speed, seen), patchSpeedSpecial)) {
                    return false;
                }
            }
            break;
        }
        const int nextLane = SB2_CHK(c.link_via[link] >= 0 ? c.link_via[link] : c.link_to[link], 0, c.nL, 36);
        const uint8_t state = c.link_state[link];
        if (link_opened(c, link, c.cur, arrivalTime, speed, speed, cf.length(), veh_impatience(c, v), cf.decel(), 0, false, v) == OPEN_NO
                || !state_priority(state)
                || state == 'Z') {
            const double remaining = seen - 0.;
            if (ins_checkFailure(cf, speed, dist, cf_insertionStopSpeed(cf, speed, remaining), patchSpeedSpecial)) {
                return false;
            }
            if (seen >= cf.minGap()) {
                break;
            }
        } else if (lane_internal(c, nextLane)) {
            double tmp = 0;
            bool dummyReq = true;
            double nSpeed = speed;
            veh_checkLinkLeader(c, e, c.lane_link_begin[nextLane], nextLane, seen + c.lane_length[nextLane], nullptr, nSpeed, tmp, tmp, dummyReq, c.cur, teleport);
            if (ins_checkFailure(cf, speed, dist, nSpeed, patchSpeed)) {
                return false;
            }
        }
        // leader on next lane: getLastVehicleInformation(aVehicle, 0, 0)
        const int nextLeader = SB2_CHK(ins_lane_last_any(c, nextLane, !teleport), -1, c.nV, 37);
        if (nextLeader >= 0 && nextLeader != v) {
            const double nextLaneSpeed = lane_safeInsertionSpeed(c, nextLane, cf, seen, nextLeader, speed);
            if (nextLaneSpeed == SB2_INVALID_SPEED || ins_checkFailure(cf, speed, dist, nextLaneSpeed, patchSpeed)) {
                return false;
            }
        }
        const double freeSpeed = cf_freeSpeed(cf, speed, seen, lane_vehicle_max_speed(c, nextLane, v), true);
        if (freeSpeed < speed) {
            if (patchSpeedSpecial) {
                speed = freeSpeed;
                dist = cf_brakeGap(cf, speed) + cf.minGap();
            }
            // else: warning only (gCheckRoutes false is not the default; the reference deschedules -> treat as failure)
            else {
                return false;
            }
        }
        const SUMOTime leaveTime = link_getLeaveTime(c, link, arrivalTime, speed, speed, cf.length());
        if (link_hasApproachingFoe(c, link, c.cur, arrivalTime, leaveTime, speed, cf.decel())) {
            if (ins_checkFailure(cf, speed, dist, cf_insertionStopSpeed(cf, speed, seen), patchSpeed)) {
                return false;
            }
        }
        arrivalTime += TIME2STEPS(c.lane_length[nextLane] / MAX2(speed, SB2_NUMERICAL_EPS));
        seen += c.lane_length[nextLane];
        currentLane = nextLane;
        if (c.link_via[link] < 0) {
            nRouteSuccs++;
        }
    }
    // followers
    const double backOffset = pos - cf.length();
    const FollowerInfo fi = lane_followers_on_consecutive(c, l, v, cf.length(), cf.decel(), backOffset, egoSpeed, !teleport);
    if (SB2_CHK(fi.veh, -1, c.nV, 41) >= 0) {
        const CF fcf = make_cf(c, c.veh_type[fi.veh]);
        const double backGapNeeded = cf_getSecureGap(fcf, c.veh_speed[fi.veh], speed, cf.decel());
        if (fi.dist < backGapNeeded) {
            return false;
        }
    }
    if (speed < 0) {
        return false;
    }
    // ---- incorporateVehicle + enterLaneAtInsertion: the vehicle's state first, then (after a fence) the queue links, so that a
    // thread of another lane that meets the new entry finds it complete (and skips it, see ins_is_new)
    if (!teleport) {
        c.veh_depart_time[v] = c.t;
        c.veh_routepos[v] = 0;
        c.veh_prevspeed[v] = speed;
        c.veh_waiting[v] = 0;
        c.veh_since_startup[v] = 86400000;   // MSVehicle ctor: myTimeSinceStartup(TIME2STEPS(3600 * 24)), MSVehicle.cpp:1023
        c.veh_jentry[v] = SB2_TIME_MAX; c.veh_jentry_ny[v] = SB2_TIME_MAX; c.veh_jconflict[v] = SB2_TIME_MAX;
        // MSDevice_Tripinfo::notifyEnter(NOTIFICATION_DEPARTED)
        c.veh_trip_routelen[v] = -pos;
        c.veh_trip_departpos[v] = pos;
        c.veh_trip_departspeed[v] = speed;
        c.veh_trip_wait[v] = 0; c.veh_trip_waitcount[v] = 0; c.veh_trip_amwaiting[v] = 0;
        c.veh_arrival_time[v] = -1; c.veh_arrival_lane[v] = -1;
        c.veh_odometer[v] = 0;
        c.veh_timeloss[v] = 0;
        veh_e1_add_lane(c, v, l, true, pos, pos - cf.length());   // addReminder + activateReminders(NOTIFICATION_DEPARTED)
        veh_e2_add_lane(c, v, l, true, pos - cf.length());
        md_count(c, l, MD_DEPARTED);
    } else {
        // MSVehicle::enterLaneAtInsertion(NOTIFICATION_TELEPORT) (MSVehicle.cpp:5793-5858): "avoid startup-effects after teleport"
        c.veh_since_startup[v] = cf.startupDelay() + c.DELTA_T;
    }
    c.veh_lane[v] = l;
    c.veh_qlane[v] = l;
    c.veh_pos[v] = pos;
    c.veh_speed[v] = speed;
    c.veh_backpos[v] = pos - cf.length();
    c.veh_accel[v] = 0;
    c.veh_lastdist[v] = 0;
    if (!teleport) {
        c.veh_havetowait[v] = 0;   // (myHaveToWaitOnNextLink is a member the constructor clears: a vehicle that ends a teleport keeps the
    }                              // value of its last executeMove, and checkRewindLinkLanes of its followers reads it, MSVehicle.cpp:5356)
    c.veh_moved[v] = 0;
    c.veh_nfurther[v] = 0;
    c.veh_stopspeed[v] = SB2_DBL_MAX;
    // position in the queue: before the first vehicle (from the rear) with pos >= ours
    int behind = -1;              // vehicle that ends up directly behind the new one
    int aheadV = c.lane_last[l];  // walks rear -> front
    while (aheadV >= 0 && !(c.veh_pos[aheadV] >= pos)) {
        behind = aheadV;
        aheadV = c.veh_ahead[aheadV];
    }
    c.veh_ahead[v] = aheadV;
    c.veh_behind[v] = behind;
#if defined(__CUDA_ARCH__)
    __threadfence();
#endif
    if (aheadV >= 0) {
        c.veh_behind[aheadV] = v;
    } else {
        c.lane_first[l] = v;
    }
    if (behind >= 0) {
        c.veh_ahead[behind] = v;
    } else {
        c.lane_last[l] = v;
    }
    c.lane_nveh[l] += 1;
    c.lane_brutto[l] += cf.length() + cf.minGap();
    c.lane_changed[l] = 1;   // MSLane::incorporateVehicle -> MSEdgeControl::gotActive (MSLane.cpp:471)
    c.edge_delayed[c.lane_edge[l]] = 1;   // MSEdge::markDelayed (MSLane.cpp:469)
    if (!teleport) {
        // MSVehicleControl::vehicleDeparted
        atomic_max_double(&c.fleet_next[0], c.veh_speedfactor[v]);
        atomic_min_double(&c.fleet_next[1], cf.decel());
        atomic_max_double(&c.fleet_next[2], cf.minGap());
        count_event(c, l, 2);
    }
    return true;
}

// MSLane::insertVehicle (departure): depart speed and position, then isInsertionSuccess; returns true if the vehicle was inserted
SB2_HD bool lane_tryInsert(const Ctx& c, int l, int v) {
    CF cf = make_cf(c, c.veh_type[v]);
    cf.desSpeed = lane_vehicle_max_speed(c, l, v);   // setTentativeLaneAndPosition (MSLane.cpp:851): veh->getLane() == this lane
    // ---- MSLane::getDepartSpeed
    double speed = 0;
    bool patchSpeed = true;
    const int dsd = c.veh_departspeed_proc[v];
    switch (dsd) {
        case DSD_GIVEN: speed = c.veh_departspeed[v]; patchSpeed = false; break;
        case DSD_MAX: speed = lane_vehicle_max_speed(c, l, v); patchSpeed = true; break;
        case DSD_DESIRED: speed = lane_vehicle_max_speed(c, l, v); patchSpeed = false; break;
        case DSD_LIMIT: speed = lane_vehicle_max_speed(c, l, v) / c.veh_speedfactor[v]; patchSpeed = false; break;
        case DSD_LAST: {
            const int last = lane_last_any(c, l);
            speed = lane_vehicle_max_speed(c, l, v);
            if (last >= 0) {
                speed = MIN2(speed, c.veh_speed[last]);
                patchSpeed = false;
            }
            break;
        }
        case DSD_AVG: {
            // MSLane::getMeanSpeed (MSLane.cpp:3477-3494): summed rear to front
            double mean = c.lane_speed[l];
            if (c.lane_nveh[l] > 0) {
                double sum = 0;
                int num = 0;
                for (int x = c.lane_last[l]; x >= 0; x = c.veh_ahead[x]) {
                    sum += c.veh_speed[x];
                    num++;
                }
                mean = sum / num;
            }
            speed = MIN2(lane_vehicle_max_speed(c, l, v), mean);
            if (lane_last_any(c, l) >= 0) {
                patchSpeed = false;
            }
            break;
        }
        default: speed = 0; patchSpeed = false; break;
    }
    // ---- MSLane::insertVehicle: position
    double pos;
    if (c.veh_departpos_proc[v] == DPD_GIVEN) {
        pos = c.veh_departpos[v];
        if (pos < 0.) {
            pos += c.lane_length[l];
        }
    } else {
        pos = MIN2(cf.length() + SB2_POSITION_EPS, c.lane_length[l]);   // basePos (edge min length == lane length: one lane per edge)
    }
    if (pos < 0 || pos > c.lane_length[l]) {
        pos = c.lane_length[l];
    }
    const bool success = lane_isInsertionSuccess(c, l, v, speed, pos, patchSpeed, patchSpeed || dsd != DSD_GIVEN, false);
    if (success && c.extrapolateDepartPos && c.t - c.veh_depart_param[v] > 0) {   // (veh_depart is rounded up to the step)
        // --extrapolate-departpos (MSLane::insertVehicle, MSLane.cpp:797-815): the part of the step the vehicle is late is made up for by
        // moving it forward at once, as far as its leader allows (MSLane::getLeader, MSLane.cpp:2957-3025; MSVehicle::executeFractionalMove)
        const SUMOTime relevantDelay = c.t - c.veh_depart_param[v] < c.DELTA_T ? c.t - c.veh_depart_param[v] : c.DELTA_T;
        const double vpos = c.veh_pos[v];
        const double vspeed = c.veh_speed[v];
        double dist = vspeed * STEPS2TIME(relevantDelay);
        int leader = -1;
        {
            AnyVehIter it = any_begin(c, l);
            for (int veh = any_next(c, l, it); veh >= 0; veh = any_next(c, l, it)) {
                if (veh != v && veh_pos_on_lane(c, veh, l) >= vpos) {
                    leader = veh;
                    break;
                }
            }
        }
        if (leader >= 0) {
            const double gap = veh_back_pos_on_lane(c, leader, l) - cf.minGap() - vpos;
            const CF lcf = make_cf(c, c.veh_type[leader]);
            const double frontGapNeeded = cf_getSecureGap(cf, vspeed, c.veh_speed[leader], lcf.decel());
            dist = MIN2(dist, gap - frontGapNeeded);
        } else if (c.lane_length[l] - vpos <= cf_brakeGap(cf, vspeed) + cf.minGap()) {
            raise(c, ERR_UNSUPPORTED_LINKSTATE);   // the leader would have to be searched beyond the lane's end (getLeaderOnConsecutive): not modelled
        }
        if (dist > 0) {
            if (vpos + dist > c.lane_length[l]) {
                raise(c, ERR_UNSUPPORTED_LINKSTATE);   // executeFractionalMove across a lane end: not modelled
            } else {
                c.veh_pos[v] = vpos + dist;
                c.veh_lastdist[v] = dist;
                c.veh_backpos[v] = vpos + dist - cf.length();
            }
        }
    }
    return success;
}

// all due candidates of departure lane l, in load order
SB2_HDN void lane_insertVehicles(const Ctx& c, int l) {
    int head = c.dq_head[l];
    const int end = c.dq_begin[l + 1];
    while (head < end) {
        const int v = SB2_CHK(c.dq_veh[SB2_CHK(head, 0, c.nV, 30)], 0, c.nV, 31);
        if (c.veh_depart[v] > c.t) {
            break;
        }
        if (c.veh_lane[v] == -4) {
            head++;   // discarded earlier while it was not at the head of the queue
            continue;
        }
        if (!lane_tryInsert(c, l, v)) {
            if (c.maxDepartDelay >= 0) {
                // MSInsertionControl::tryInsert (MSInsertionControl.cpp:170-173): every pending vehicle of this edge fails in this step
                // (the edge remembers the failed attempt, MSEdge::insertVehicle); the ones that have waited too long are deleted
                for (int k = head; k < end; k++) {
                    const int w = c.dq_veh[k];
                    if (c.veh_depart[w] > c.t) {
                        break;
                    }
                    if (c.veh_lane[w] == -1 && c.t - c.veh_depart_param[w] > c.maxDepartDelay) {
                        c.veh_lane[w] = -4;
                        atomic_inc(&c.counters[15]);
                    }
                }
                while (head < end && c.veh_lane[c.dq_veh[head]] == -4) {
                    head++;
                }
            }
            break;
        }
        head++;
    }
    c.dq_head[l] = head;
}

}  // namespace sb2
