/* sb2_link.h -- right-of-way at junctions: the approach table and MSLink::opened / blockedAtTime / blockedByFoe.
 *
 * Restates src/microsim/MSLink.cpp:829-833 (getLeaveTime), 837-1022 (opened; non-sublane, no bidi, no persons),
 * 1026-1059 (blockedAtTime), 1113-1221 (blockedByFoe), 1225-1285 (computeFoeArrivalTimeBraking), 1289-1301
 * (hasApproachingFoe), 1346-1397 (isCont, lastWasContMajor, lastWasContState), MSLink.h:730 (unsafeMergeSpeeds).
 *
 * The approach table of the reference (std::map<vehicle, ApproachingVehicleInformation> per link, MSLink.h:183) is
 * the set of drive items of one buffer, chained per link through AppRec::next (lock-free push in k_plan).  Iteration order
 * differs from the reference's map order, which is unobservable here: blockedAtTime is a pure OR over the foes unless
 * collectFoes is used (zipper links only; refused at load).
 */
#pragma once
#include "sb2_types.h"

namespace sb2 {

SB2_HD bool veh_on_road(const Ctx& c, int v) { return c.veh_lane[v] >= 0; }

// MSLink.cpp:1346-1349
SB2_HD bool link_isCont(const Ctx& c, int k) {
    const uint8_t s = c.link_state[k];
    return (s == 'o' || s == 'w' || s == 's') ? (c.link_flags[k] & KF_AMCONT_OFF) != 0 : (c.link_flags[k] & KF_AMCONT) != 0;
}

// MSLink.cpp:1353-1378
SB2_HD bool link_lastWasContMajor(const Ctx& c, int k) {
    if (c.link_flags[k] & KF_EXIT_AFTER_IJ) {
        k = c.inc_link[c.lane_inc_begin[c.link_internal_before[k]]];
    }
    if (c.link_via[k] < 0 || (c.link_flags[k] & KF_AMCONT)) {
        return false;
    }
    const int predLink = c.link_contpred[k];
    if (predLink < 0) {
        return false;
    }
    if (state_priority(c.link_state[predLink])) {
        return true;
    }
    if (c.link_flags[k] & KF_PEDFOE) {
        return c.link_last_green[predLink] == 'G';
    }
    return state_yellow(c.link_state[predLink]);
}

// MSLink.cpp:1382-1397
SB2_HD bool link_lastWasContState(const Ctx& c, int k, uint8_t linkState) {
    if (c.link_via[k] < 0 || (c.link_flags[k] & (KF_AMCONT | KF_PEDFOE))) {
        return false;
    }
    const int predLink = c.link_contpred[k];
    if (predLink < 0) {
        return false;
    }
    return c.link_state[predLink] == linkState;
}

// MSLink.cpp:829-833
SB2_HD SUMOTime link_getLeaveTime(const Ctx& c, int k, SUMOTime arrivalTime, double arrivalSpeed, double leaveSpeed, double vehicleLength) {
    return arrivalTime == SB2_TIME_MAX ? SB2_TIME_MAX
           : arrivalTime + TIME2STEPS((c.link_length[k] + vehicleLength) / MAX2(0.5 * (arrivalSpeed + leaveSpeed), SB2_NUMERICAL_EPS));
}

// MSLink.h:730
SB2_HD bool unsafeMergeSpeeds(double leaderSpeed, double followerSpeed, double leaderDecel, double followerDecel) {
    return div_pos(leaderSpeed * leaderSpeed, leaderDecel) <= div_pos(followerSpeed * followerSpeed, followerDecel);   // decels > 0
}

// MSLink.cpp:1225-1285
SB2_HD SUMOTime computeFoeArrivalTimeBraking(const Ctx& c, SUMOTime arrivalTime, int foe, SUMOTime foeArrivalTime, double impatience, double dist, double& fasb) {
    if (arrivalTime - arrivalTime % c.DELTA_T == foeArrivalTime - foeArrivalTime % c.DELTA_T) {
        return foeArrivalTime;
    }
    if (arrivalTime % c.DELTA_T > 0) {
        arrivalTime = arrivalTime - (arrivalTime % c.DELTA_T) + c.DELTA_T;
    }
    const double m = c.vt_decel[c.veh_type[foe]] * impatience;
    const double dt = STEPS2TIME(foeArrivalTime - arrivalTime);
    const double d = dt * m;
    const double a = dt * d / 2;
    const double v = dist / STEPS2TIME(foeArrivalTime - c.t + c.DELTA_T);
    const double dist2 = dist - v * STEPS2TIME(arrivalTime - c.t);
    if (0.5 * v * v / m <= dist2) {
        fasb = 0;
        return foeArrivalTime + TIME2STEPS(30);
    }
    const double x = (sqrt(4 * (v - d) * (v - d) - 8 * m * a) * -0.5 - d + v) / m;
    fasb = v - (dt + x) * m;
    return foeArrivalTime + TIME2STEPS(x);
}

// MSLink.cpp:1113-1221 ; `k` is the foe link whose approach record `r` (of vehicle `foe`) is tested
SB2_HD bool link_blockedByFoe(const Ctx& c, int k, int buf, int r, const AppRec& a,
                              SUMOTime arrivalTime, SUMOTime leaveTime, double arrivalSpeed, double leaveSpeed,
                              bool sameTargetLane, double impatience, double decel, SUMOTime waitingTime) {
    if (a.veh < 0) {   // !avi.willPass
        return false;
    }
    const int foe = a.veh;
    const uint8_t state = c.link_state[k];
    const SUMOTime aviArrival = a.arrtime;
    const SUMOTime aviLeaving = a.leavetime;
    if (state == 'w') {
        // action step == sim step: actionDelta = SIMSTEP - lastActionTime = 0 during executeMove
        const SUMOTime actionDelta = 0;
        if (waitingTime > c.di_regwaiting[buf][r] + actionDelta) {
            return false;
        }
        if (waitingTime == (c.di_regwaiting[buf][r] + actionDelta) && arrivalTime < aviArrival + actionDelta) {
            return false;
        }
    }
    SUMOTime foeArrivalTime = aviArrival;
    double foeArrivalSpeedBraking = a.arrspeedbraking;
    if (impatience > 0 && arrivalTime < aviArrival) {
        const SUMOTime fatb = computeFoeArrivalTimeBraking(c, arrivalTime, foe, aviArrival, impatience, c.di_dist[buf][r], foeArrivalSpeedBraking);
        foeArrivalTime = (SUMOTime)((1. - impatience) * (double)aviArrival + impatience * (double)fatb);
    }
    const SUMOTime lookAhead = (state == 'Z') ? 16000 : 1000;   // MSLink.cpp:72-74 ; jmTimegapMinor default = myLookaheadTime
    if (aviLeaving < arrivalTime) {
        // ego wants to be follower
        const double aviLeaveSpeed = c.di_accelv[buf][r] < 0 ? c.di_vpass[buf][r] : c.di_accelv[buf][r];
        if (sameTargetLane && (arrivalTime - aviLeaving < lookAhead
                               || unsafeMergeSpeeds(aviLeaveSpeed, arrivalSpeed, c.vt_decel[c.veh_type[foe]], decel))) {
            return true;
        }
    } else if (foeArrivalTime > leaveTime + lookAhead) {
        // ego wants to be leader
        if (sameTargetLane && unsafeMergeSpeeds(leaveSpeed, foeArrivalSpeedBraking, decel, c.vt_decel[c.veh_type[foe]])) {
            return true;
        }
    } else {
        return true;
    }
    return false;
}

// how MSLink::blockedAtTime reads a foe's speed under lastWasContRed (`it.first->getSpeed()`, MSLink.cpp:1050): the reference
// executes lanes one after the other, so the speed is the foe's NEW one if the foe moved earlier in this step:
//   LIVE_NOW   the caller runs while no vehicle moves (planMove: previous-step table and start-of-step speeds; insertion: this
//              step's table and end-of-step speeds) -> veh_speed
//   LIVE_PASS1 executeMove, first pass: a foe that moves later counts with its registered (start-of-step) speed; a blocking foe
//              that moves earlier makes the answer OPEN_UNDETERMINED (the vehicle is deferred to pass 2)
//   LIVE_PASS2 executeMove, second pass: earlier foes have their new speed in veh_speed unless they are deferred themselves
enum { LIVE_NOW = 0, LIVE_PASS1 = 1, LIVE_PASS2 = 2 };
enum { OPEN_NO = 0, OPEN_YES = 1, OPEN_UNDETERMINED = 2 };

// true if the reference moves `foe` before `ego` in this step: MSEdgeControl::executeMovements walks myActiveLanes in list order
// (lane_actkey), MSLane::executeMovements walks a lane front to back
SB2_HD bool veh_executes_before(const Ctx& c, int foe, int ego) {
    const int lf = c.veh_qlane[foe], le = c.veh_qlane[ego];
    if (lf == le) {
        return c.veh_len_in_front[foe] < c.veh_len_in_front[ego];
    }
    return c.lane_actkey[lf] < c.lane_actkey[le];
}

// MSLink.cpp:1026-1059 (no persons, jmIgnoreFoeProb == 0, no ignoreFoe lists); returns OPEN_NO (= blocked) / OPEN_YES / OPEN_UNDETERMINED
SB2_HD int link_blockedAtTime(const Ctx& c, int k, int buf, SUMOTime arrivalTime, SUMOTime leaveTime, double arrivalSpeed, double leaveSpeed,
                              bool sameTargetLane, double impatience, double decel, SUMOTime waitingTime, int ego, bool lastWasContRed, int live = LIVE_NOW) {
    int result = OPEN_YES;
    for (int r = c.link_app_head[buf][k]; r >= 0;) {
        const AppRec a = c.ap_rec[buf][SB2_CHK(r, 0, (long long)c.nV * c.max_items, 8)];
        const int rec = r;
        r = a.next;
        // The reference tests, per approaching vehicle: not ego, (under lastWasContRed) not standing, blockedByFoe.  The tests are
        // independent, so the cheapest comes first: most records are dismissed from their 32-byte hot part alone (no request, or
        // their time window does not meet ego's), before anything of the vehicle itself is loaded.
        if (a.veh < 0 || a.veh == ego) {
            continue;
        }
        if (!link_blockedByFoe(c, k, buf, rec, a, arrivalTime, leaveTime, arrivalSpeed, leaveSpeed, sameTargetLane, impatience, decel, waitingTime)) {
            continue;
        }
        const int foe = SB2_CHK(a.veh, 0, c.nV, 9);
        if (!veh_on_road(c, foe)) {
            continue;
        }
        if (lastWasContRed) {
            double foeSpeed;
            if (live == LIVE_NOW) {
                foeSpeed = c.veh_speed[foe];
            } else if (!veh_executes_before(c, foe, ego)) {
                foeSpeed = c.di_regspeed[buf][rec];
            } else if (live == LIVE_PASS2 && !c.veh_deferred[foe]) {
                foeSpeed = c.veh_speed[foe];
            } else {
                result = OPEN_UNDETERMINED;   // the foe's new speed is not known yet, and this record would block
                continue;
            }
            if (!(foeSpeed > SB2_HALTING_SPEED)) {
                continue;
            }
        }
        return OPEN_NO;
    }
    return result;
}

// MSLink.cpp:837-1022; OPEN_NO / OPEN_YES, or OPEN_UNDETERMINED in executeMove's first pass (see LIVE_PASS1)
SB2_RARE int link_opened(const Ctx& c, int k, int buf, SUMOTime arrivalTime, double arrivalSpeed, double leaveSpeed, double vehicleLength,
                        double impatience, double decel, SUMOTime waitingTime, bool ignoreRed, int ego, int live = LIVE_NOW) {
    const uint8_t state = c.link_state[k];
    if (state_red(state) && !ignoreRed) {
        return OPEN_NO;
    }
    if (link_isCont(c, k)) {
        return OPEN_YES;
    }
    const SUMOTime leaveTime = link_getLeaveTime(c, k, arrivalTime, arrivalSpeed, leaveSpeed, vehicleLength);
    if ((state_priority(state)
            || link_lastWasContState(c, k, 'G')
            || ((c.link_flags[k] & KF_EXIT_AFTER_IJ) && c.link_state[c.link_entry[k]] == 'G'))
            && state != 'Z') {
        return OPEN_YES;
    }
    // jmAllwayStopWait / jmStopSignWait default TS
    if (state == 'w' && waitingTime < TIME2STEPS(c.TS)) {
        return OPEN_NO;
    } else if (state == 's' && waitingTime < TIME2STEPS(c.TS)) {
        return OPEN_NO;
    }
    const bool lastWasContRed = link_lastWasContState(c, k, 'r');
    const int to = c.link_to[k];
    int result = OPEN_YES;
    for (int i = c.link_foelink_begin[k]; i < c.link_foelink_begin[k + 1]; i++) {
        const int fl = c.foelink[i];
        const int b = link_blockedAtTime(c, fl, buf, arrivalTime, leaveTime, arrivalSpeed, leaveSpeed, to == c.link_to[fl],
                                         impatience, decel, waitingTime, ego, lastWasContRed, live);
        if (b == OPEN_NO) {
            return OPEN_NO;
        }
        if (b == OPEN_UNDETERMINED) {
            result = OPEN_UNDETERMINED;
        }
    }
    return result;
}

SB2_HD int lane_nveh_with_partials(const Ctx& c, int l) { return c.lane_nveh[l] + c.lane_npartial[l]; }

// MSLink.cpp:1289-1301
SB2_HD bool link_hasApproachingFoe(const Ctx& c, int k, int buf, SUMOTime arrivalTime, SUMOTime leaveTime, double speed, double decel) {
    const int to = c.link_to[k];
    for (int i = c.link_foelink_begin[k]; i < c.link_foelink_begin[k + 1]; i++) {
        const int fl = c.foelink[i];
        if (link_blockedAtTime(c, fl, buf, arrivalTime, leaveTime, speed, speed, to == c.link_to[fl], 0, decel, 0, -1, false) == OPEN_NO) {
            return true;
        }
    }
    for (int i = c.link_foelane_begin[k]; i < c.link_foelane_begin[k + 1]; i++) {
        if (c.foelane[i] >= 0 && lane_nveh_with_partials(c, c.foelane[i]) > 0) {
            return true;
        }
    }
    return false;
}

// find the approach record of vehicle v on link k (MSLink::getApproaching, MSLink.cpp:801-809); -1 if absent
SB2_HD int link_getApproaching(const Ctx& c, int k, int buf, int v) {
    for (int r = c.link_app_head[buf][k]; r >= 0;) {
        const AppRec a = c.ap_rec[buf][r];
        if (a.veh == v || a.veh == ~v) {
            return r;
        }
        r = a.next;
    }
    return -1;
}

}  // namespace sb2
