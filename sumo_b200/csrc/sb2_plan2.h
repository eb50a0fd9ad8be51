/* sb2_plan2.h -- planMoveInternal, checkRewindLinkLanes and approach registration for one vehicle.
 * Reference lines: see sb2_plan.h.
 */
#pragma once
#include "sb2_plan.h"

namespace sb2 {

// the drive items of the vehicle at work-list position wp, as they lie in the item-major SoA of buffer c.cur (item i at i * nV + wp:
// a warp's accesses to "item i" are coalesced).  planMove writes every item ONCE when the look-ahead has finished with it and
// checkRewindLinkLanes edits them in place -- there is no per-thread DriveItem array (it was 640 B of local memory per thread,
// most of the register-spill traffic of the kernel: profiles/r01_k_plan_ncu.txt)
struct DIView {
    const Ctx& c;
    int buf;
    size_t wp, nV;
    SB2_HD size_t at(int i) const { return (size_t)i * nV + wp; }
    SB2_HD int link(int i) const { return c.di_link[buf][at(i)]; }
    SB2_HD bool request(int i) const { return c.di_request[buf][at(i)] != 0; }
    SB2_HD double leaveSpeed(int i) const { const double a = c.di_accelv[buf][at(i)]; return a < 0 ? c.di_vpass[buf][at(i)] : a; }
    SB2_HD void store(int i, const DriveItem& d) const {
        const size_t r = at(i);
        c.di_link[buf][r] = d.link;
        c.di_vpass[buf][r] = d.vPass;
        c.di_vwait[buf][r] = d.vWait;
        c.di_request[buf][r] = d.setRequest;
        c.di_arrtime[buf][r] = d.arrivalTime;
        c.di_arrspeed[buf][r] = d.arrivalSpeed;
        c.di_arrspeedbraking[buf][r] = d.arrivalSpeedBraking;
        c.di_dist[buf][r] = d.distance;
        c.di_accelv[buf][r] = d.accelV;
        c.di_hadstopped[buf][r] = d.hadStoppedVehicle;
        c.di_avail[buf][r] = d.availableSpace;
    }
};

// MSVehicle::checkRewindLinkLanes (MSVehicle.cpp:5255-5489); tableBuf = approach table of the previous step
SB2_HD void veh_checkRewindLinkLanes(const Ctx& c, const Ego& e, double lengthsInFront, const DIView& lf, int n, int tableBuf) {
    if (c.lane_flags[e.lane] & LF_ROUNDABOUT) {
        return;
    }
    const int cur = lf.buf;
    const double lengthWithGap = e.length + e.minGap;
    double seenSpace = -lengthsInFront;
    bool foundStopped = false;
    for (int i = 0; i < n; ++i) {
        const size_t r = lf.at(i);
        const int link = c.di_link[cur][r];
        if (link < 0 || foundStopped) {
            if (!foundStopped) {
                c.di_avail[cur][r] += seenSpace;
            } else {
                c.di_avail[cur][r] = seenSpace;
            }
            continue;
        }
        int approachedLane = c.link_via[link];
        if (approachedLane >= 0) {
            if (veh_keepClear(c, link)) {
                seenSpace = seenSpace - c.lane_brutto[approachedLane];
                if (approachedLane == e.lane) {
                    seenSpace += lengthWithGap;
                }
            } else {
                if (c.ls_found_stopped[approachedLane]) {
                    foundStopped = true;
                }
                seenSpace = seenSpace + c.ls_space[approachedLane];
            }
            c.di_avail[cur][r] = seenSpace;
            continue;
        }
        approachedLane = c.link_to[link];
        const int last = lane_last_any(c, approachedLane);
        if (last < 0 || last == e.v) {
            if (c.lane_length[approachedLane] > e.length || veh_keepClear(c, link)) {
                seenSpace += c.lane_length[approachedLane];
            }
            c.di_avail[cur][r] = seenSpace;
        } else {
            const bool foundStopped2 = c.ls_found_stopped[approachedLane] != 0;
            const double spaceTillLastStanding = c.ls_space[approachedLane];
            seenSpace += spaceTillLastStanding;
            if (foundStopped2) {
                foundStopped = true;
                c.di_hadstopped[cur][r] = 1;
            }
            c.di_avail[cur][r] = seenSpace;
            if (c.veh_havetowait[last]) {
                foundStopped = true;
                c.di_hadstopped[cur][r] = 1;
            }
        }
    }
    // check which links allow continuation and add pass available to the previous item
    for (int i = n - 1; i > 0; --i) {
        const size_t r = lf.at(i - 1), rNext = lf.at(i);
        const int itemLink = c.di_link[cur][r];
        const int nextLink = c.di_link[cur][rNext];
        // (the reference dereferences item.myLink before testing it; a link-less item can only be the last one)
        const bool canLeaveJunction = itemLink < 0 || c.link_via[itemLink] < 0 || nextLink < 0 || c.di_request[cur][rNext] != 0;
        const bool opened = (itemLink >= 0
                             && (canLeaveJunction || (
                                     nextLink >= 0 && (c.link_flags[nextLink] & KF_INTERNAL_JUNCTION) && state_red(c.link_state[nextLink])))
                             && (
                                 state_priority(c.link_state[itemLink])
                                 || i == 1
                                 || link_opened(c, itemLink, tableBuf, c.di_arrtime[cur][r], c.di_arrspeed[cur][r],
                                                lf.leaveSpeed(i - 1), e.length,
                                                veh_impatience(c, e.v), e.cf.decel(), e.waiting, false, e.v) != OPEN_NO));
        bool allowsContinuation = (itemLink < 0 || link_isCont(c, itemLink) || opened) && !c.di_hadstopped[cur][r];
        if (!opened && itemLink >= 0) {
            foundStopped = true;
            if (i > 1) {
                const int link2 = c.di_link[cur][lf.at(i - 2)];
                if (link2 >= 0 && link_isCont(c, link2)) {
                    allowsContinuation = true;
                }
            }
        }
        if (allowsContinuation) {
            c.di_avail[cur][r] = c.di_avail[cur][rNext];
        }
    }
    // find removalBegin
    int removalBegin = -1;
    for (int i = 0; foundStopped && i < n && removalBegin < 0; ++i) {
        const size_t r = lf.at(i);
        const int link = c.di_link[cur][r];
        if (link < 0) {
            continue;
        }
        const double leftSpace = c.di_avail[cur][r] - lengthWithGap;
        if (leftSpace < 0) {
            double impatienceCorrection = 0;
            if (leftSpace < -impatienceCorrection / 10. && veh_keepClear(c, link)) {
                removalBegin = i;
            }
        }
    }
    // abort requests
    if (removalBegin != -1 && !(removalBegin == 0 && lane_internal(c, e.lane))) {
        const double brakeGap = cf_brakeGap(e.cf, e.speed, e.cf.decel(), 0.);
        while (removalBegin < n) {
            const size_t r = lf.at(removalBegin);
            const int link = c.di_link[cur][r];
            if (link < 0) {
                break;
            }
            c.di_vpass[cur][r] = c.di_vwait[cur][r];
            if (c.di_dist[cur][r] >= brakeGap + SB2_POSITION_EPS) {
                if (!(c.link_flags[link] & KF_EXIT) || !c.di_request[cur][lf.at(removalBegin - 1)]) {
                    c.di_request[cur][r] = 0;
                }
            }
            ++removalBegin;
        }
    }
}

// MSVehicle::planMove + planMoveInternal for vehicle v at time t; writes the drive items of buffer c.cur and registers
// them in the approach table of buffer c.cur (MSVehicle::setApproachingForAllLinks).
// wp = position of v in this step's work list (addresses its drive-item records)
SB2_HDN void veh_planMove(const Ctx& c, int v, int wp) {
    (void)SB2_CHK(v, 0, c.nV, 2);
    const int myLane = SB2_CHK(c.veh_lane[v], -2, c.nL, 3);
    if (myLane < 0) {
        return;
    }
    const SUMOTime t = c.t;
    const int cur = c.cur, prev = c.cur ^ 1;
    Ego e;
    e.v = v; e.vt = c.veh_type[v]; e.lane = myLane;
    e.pos = c.veh_pos[v]; e.speed = c.veh_speed[v];
    e.cf = make_cf(c, e.vt);
    e.cf.desSpeed = lane_vehicle_max_speed(c, myLane, v);
    e.length = e.cf.length(); e.minGap = e.cf.minGap();
    e.waiting = c.veh_waiting[v];
    const CF& cf = e.cf;

    // MSVehicle.cpp:2183
    c.veh_stopspeed[v] = cf_maxNextSpeed(cf, c.veh_stopspeed[v]);

    // items 0 .. n-2 are in the drive-item arrays already; item n-1 (`last`, the only one the look-ahead still edits through
    // lastLink) stays in registers until the next one is pushed
    const DIView lf = {c, cur, (size_t)wp, (size_t)c.nV};
    DriveItem last;
    int n = 0;
    bool overflow = false;
#define SB2_PUSH(item) do { if (n < c.max_items) { if (n > 0) { lf.store(n - 1, last); } last = (item); n++; } else { overflow = true; } } while (0)

    // leader on the own lane (MSLane::planMovements / updateLeaderInfo, MSLane.cpp:1608-1685)
    int ahead = c.veh_ahead[v];
    if (ahead < 0 && c.lane_npartial[myLane] > 0) {
        ahead = SB2_CHK(c.lane_partial[(size_t)myLane * MAX_PARTIAL], 0, c.nV, 1);
    }

    const double vehicleLength = e.length;
    const double maxV = cf_maxNextSpeed(cf, e.speed);
    const double vehMaxSpeed = veh_max_speed(c, v);
    const double maxVD = MAX2(vehMaxSpeed, MIN2(maxV, veh_desired_max_speed(c, v)));
    double laneMaxV = lane_vehicle_max_speed(c, myLane, v, maxVD);
    const double vMinComfortable = cf_minNextSpeed(cf, e.speed);
    laneMaxV = MAX2(laneMaxV, vMinComfortable);
    // maximumLaneSpeedCF (MSCFModel.h:245)
    double vv = MIN2(maxV, laneMaxV);
    if (c.traci_speed_used && c.veh_traci_speed[v] >= 0.) {
        // MSVehicle.cpp:2313-2326: v = myInfluencer->influenceSpeed(t, v, v, MAX2(0, minNextSpeed), maxV)
        vv = traci_influenceSpeed(c.veh_traci_speed[v], vv, MAX2(0., vMinComfortable), maxV);
    }
    const double dist = maxV * c.TS + cf_brakeGap(cf, maxV);
    bool hadNonInternal = false;
    double seen = c.lane_length[myLane] - e.pos;
    double seenNonInternal = 0;
    double seenInternal = lane_internal(c, myLane) ? seen : 0;
    double vLinkPass = MIN2(cf_estimateSpeedAfterDistance(cf, seen, vv, cf.accel()), laneMaxV);
    int view = 0;
    DriveItem* lastLink = nullptr;
    bool slowedDownForMinor = false;
    int lane = myLane;
    int leaderLane = myLane;
    const int routeLen = veh_route_len(c, v);
    const int routePos = c.veh_routepos[v];
    while (true) {
        veh_adaptToLeaders(c, e, ahead, seen, lastLink, leaderLane, vv, vLinkPass);
        if (lastLink != nullptr) {
            lastLink->vWait = MIN2(lastLink->vWait, vv);
        }
        // move to next lane
        const int link = lane_succ_link(c, lane, veh_succ_edge(c, v, view + 1));
        // check whether the vehicle is on its final edge
        if (routePos + view + 1 == routeLen) {
            const double arrivalSpeed = c.veh_arrivalspeed[v] >= 0. ? c.veh_arrivalspeed[v] : laneMaxV;   // ArrivalSpeedDefinition::GIVEN
            const double distToArrival = seen + c.veh_arrivalpos[v] - c.lane_length[lane] - arrivalSpeed * c.TS;
            const double va = MAX2(SB2_NUMERICAL_EPS, cf_freeSpeed(cf, e.speed, distToArrival, arrivalSpeed, false));
            vv = MIN2(vv, va);
            if (lastLink != nullptr) {
                di_adaptLeaveSpeed(*lastLink, va);
            }
            SB2_PUSH(di_make_nolink(vv, seen, (c.lane_flags[lane] & LF_FRINGE) ? 1000 : 0));
            break;
        }
        // dead end
        if (link < 0) {
            double va = cf_stopSpeed(cf, e.speed, seen);
            if (lastLink != nullptr) {
                di_adaptLeaveSpeed(*lastLink, va);
            }
            vv = MIN2(va, vv);
            SB2_PUSH(di_make_nolink(vv, seen, 0));
            break;
        }
        if (c.lane_vis != nullptr && c.lane_vis[myLane] == LV_AUTH && c.lane_vis[c.link_from[link]] < LV_GHOST) {
            // the look-ahead of one of the rank's OWN vehicles reaches a junction whose approaching vehicles this rank does not plan:
            // the ghost column is too narrow for this network (edges shorter than the look-ahead) -- reported, never silently wrong.
            // (Ghost vehicles heading away from the strip do look at such links; their result is replaced by the owner's anyway.)
            raise(c, ERR_PARTITION_REACH);
        }
        const uint8_t lstate = c.link_state[link];
        const bool yellowOrRed = state_red(lstate) || state_yellow(lstate);
        double laneStopOffset;
        const double majorStopOffset = MAX2(SB2_DIST_TO_STOPLINE_EXPECT_PRIORITY, 0.);
        const double stopDecel = yellowOrRed ? MAX2(MIN2(c.tlsYellowMinDecel, cf.emergencyDecel()), cf.decel()) : cf.decel();
        const double brakeDist = cf_brakeGap(cf, e.speed, stopDecel, 0);
        const bool canBrakeBeforeLaneEnd = seen >= brakeDist;
        const bool canBrakeBeforeStopLine = seen - 0 >= brakeDist;
        const double foeVis = c.link_foevis[link];
        if (yellowOrRed) {
            laneStopOffset = majorStopOffset;
        } else if (state_priority(lstate)) {
            laneStopOffset = MIN2(foeVis - SB2_POSITION_EPS, majorStopOffset);
        } else {
            // MAX2(stopOffset 0, jmStoplineCrossingGap(SAFETY_GAP 1) - distToFoePedCrossing(DBL_MAX)) = 0; jmStoplineGap(Minor) default 0
            double minorStopOffset = MAX2(0., 1. - SB2_DBL_MAX);
            minorStopOffset = MAX2(minorStopOffset, 0.);
            laneStopOffset = MIN2(foeVis - SB2_POSITION_EPS, minorStopOffset);
        }
        if (canBrakeBeforeLaneEnd) {
            laneStopOffset = MIN2(laneStopOffset, seen - brakeDist);
        }
        laneStopOffset = MAX2(SB2_POSITION_EPS, laneStopOffset);
        double stopDist = MAX2(0., seen - laneStopOffset);

        const bool leavingCurrentIntersection = lane_internal(c, myLane) && lastLink == nullptr;
        const bool abortRequestAfterMinor = slowedDownForMinor && c.link_internal_before[link] < 0;
        bool setRequest = (vv > SB2_NUMERICAL_EPS_SPEED && !abortRequestAfterMinor) || (leavingCurrentIntersection);

        double stopSpeed = cf_stopSpeed(cf, e.speed, stopDist, stopDecel, false);
        double vLinkWait = MIN2(vv, stopSpeed);

        // MSVehicle.cpp:2920-2945: a vehicle that can stop at yellow/red only looks for link leaders when it is already on the
        // intersection; everybody else always does (one call site for both cases keeps the kernel's code small)
        const bool stopAtRed = yellowOrRed && canBrakeBeforeStopLine /* && !ignoreRed && seen >= mustSeeBeforeReversal(0) */;
        if (!stopAtRed || lane_internal(c, lane)) {
            veh_checkLinkLeader(c, e, link, lane, seen, lastLink, vv, vLinkPass, vLinkWait, setRequest, prev);
        }
        if (stopAtRed) {
            const SUMOTime arrivalTime = veh_getArrivalTime(c, e, t, seen, vv, vLinkPass);
            SB2_PUSH(di_make(link, vv, vLinkWait, false, arrivalTime, vLinkWait, 0, seen, -1));
            break;
        }
        // (ignoreRed is always false for jmDriveAfterRedTime = -1: the restrict-speed-at-red branch :2948 is dead)

        if (lastLink != nullptr) {
            di_adaptLeaveSpeed(*lastLink, laneMaxV);
        }
        double arrivalSpeed = vLinkPass;
        const double visibilityDistance = foeVis;
        const bool determinedFoePresence = seen <= visibilityDistance;
        const bool couldBrakeForMinor = !state_priority(lstate) && brakeDist < seen && !link_lastWasContMajor(c, link);
        if (couldBrakeForMinor && !determinedFoePresence) {
            double maxSpeedAtVisibilityDist = cf_maximumSafeStopSpeed(cf, visibilityDistance, cf.decel(), e.speed, false, 0., false);
            double maxArrivalSpeed = cf_estimateSpeedAfterDistance(cf, visibilityDistance, maxSpeedAtVisibilityDist, cf.accel());
            arrivalSpeed = MIN2(vLinkPass, maxArrivalSpeed);
            slowedDownForMinor = true;
        } else if (lstate == '=' && e.waiting > 0) {
            // right-before-left deadlock breaker draws from the lane RNG (MSVehicle.cpp:3002-3020): refused at load
            raise(c, ERR_UNSUPPORTED_LINKSTATE);
        }
        const SUMOTime arrivalTime = veh_getArrivalTime(c, e, t, seen, vv, arrivalSpeed);
        if (couldBrakeForMinor && determinedFoePresence && (c.lane_flags[c.link_to[link]] & LF_ROUNDABOUT)) {
            const bool wasOpened = link_opened(c, link, prev, arrivalTime, arrivalSpeed, arrivalSpeed, e.length, veh_impatience(c, v),
                                               cf.decel(), e.waiting, false, v) != OPEN_NO;
            if (!wasOpened) {
                slowedDownForMinor = true;
            }
        }
        double arrivalSpeedBraking = 0;
        const double bGap = cf_brakeGap(cf, vv);
        if (seen < bGap) {
            arrivalSpeedBraking = cf_getMinimalArrivalSpeedEuler(cf, seen, vv);
            arrivalSpeedBraking = MIN2(arrivalSpeedBraking, arrivalSpeed);
        }
        const int nextLane = c.link_via[link] >= 0 ? c.link_via[link] : c.link_to[link];
        const double estimatedLeaveSpeed = MIN2(lane_vehicle_max_speed(c, nextLane, v, maxVD),
                                                cf_estimateSpeedAfterDistance(cf, c.link_length[link], arrivalSpeed, cf.accel()));
        SB2_PUSH(di_make(link, vv, vLinkWait, setRequest, arrivalTime, arrivalSpeed, arrivalSpeedBraking, seen, estimatedLeaveSpeed));
        if (c.link_via[link] < 0) {
            hadNonInternal = true;
            ++view;
        }
        if ((!setRequest || vv <= 0 || seen > dist) && hadNonInternal
                && seenNonInternal > MAX2(vehicleLength * SB2_CRLL_LOOK_AHEAD, vehicleLength + seenInternal)) {
            break;
        }
        if (overflow) {
            break;
        }
        // get the following lane
        lane = nextLane;
        laneMaxV = lane_vehicle_max_speed(c, lane, v, maxVD);
        const double va = MAX2(cf_freeSpeed(cf, e.speed, seen, laneMaxV, false), vMinComfortable - SB2_NUMERICAL_EPS);
        vv = MIN2(va, vv);
        if (lane_internal(c, lane)) {
            seenInternal += c.lane_length[lane];
        } else {
            seenNonInternal += c.lane_length[lane];
        }
        leaderLane = lane;
        ahead = SB2_CHK(lane_last_any(c, SB2_CHK(leaderLane, 0, c.nL, 4)), -1, c.nV, 5);   // getLastVehicleInformation(nullptr, 0)
        seen += c.lane_length[lane];
        vLinkPass = MIN2(cf_estimateSpeedAfterDistance(cf, c.lane_length[lane], vv, cf.accel()), laneMaxV);
        lastLink = &last;
    }
#undef SB2_PUSH
    if (n > 0) {
        lf.store(n - 1, last);
    }
    if (overflow) {
        raise(c, ERR_DRIVE_ITEM_OVERFLOW);
    }

    veh_checkRewindLinkLanes(c, e, c.veh_len_in_front[v], lf, n, prev);

    // register approaches (MSVehicle::setApproachingForAllLinks -> MSLink::setApproaching, MSLink.cpp:718-736)
    c.veh_ndi[cur][wp] = n;
    for (int i = 0; i < n; i++) {
        const size_t r = lf.at(i);
        const int link = c.di_link[cur][r];
        if (link >= 0) {
            const SUMOTime arrivalTime = c.di_arrtime[cur][r];
            c.di_regspeed[cur][r] = e.speed;
            c.di_regwaiting[cur][r] = e.waiting;
            AppRec a;
            a.veh = c.di_request[cur][r] ? v : ~v;
            a.arrtime = arrivalTime;
            a.leavetime = link_getLeaveTime(c, link, arrivalTime, c.di_arrspeed[cur][r], lf.leaveSpeed(i), e.length);
            a.arrspeedbraking = c.di_arrspeedbraking[cur][r];
            a.next = atomic_exch(&c.link_app_head[cur][link], (int32_t)r);
            c.ap_rec[cur][r] = a;
        }
    }
}

}  // namespace sb2
