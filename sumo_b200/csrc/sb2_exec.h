/* sb2_exec.h -- MSVehicle::executeMove for one vehicle (one GPU thread per vehicle).
 *
 * Restates src/microsim/MSVehicle.cpp:3744-4021 (processLinkApproaches), 4633-4905 (executeMove), 4965-5018 (updateState),
 * 4291-4315 (updateWaitingTime, updateTimeLoss), 4469-4628 (processLaneAdvances), 5055-5131 (updateFurtherLanes),
 * 5565-5624 (enterLaneAtMove), 5862-5885 (leaveLane: odometer), 1130-1135 (hasArrivedInternal);
 * src/microsim/cfmodels/MSCFModel.cpp:199-285 (finalizeSpeed, applyStartupDelay), MSCFModel_Krauss.cpp:94-123
 * (patchSpeedBeforeLC).
 *
 * Reads of other vehicles: only their approach records of this step (buffer c.cur) and static type data, so all vehicles
 * execute in parallel.  Lane membership is not touched here: a vehicle that advanced pushes itself on the incoming list
 * of its new lane (MSLane::myVehBuffer) and k_handoff rebuilds the lane queues.
 */
#pragma once
#include "sb2_types.h"
#include "sb2_cf.h"
#include "sb2_veh.h"
#include "sb2_link.h"
#include "sb2_emis.h"
#include "sb2_e1.h"
#include "sb2_e2.h"
#include "sb2_meandata.h"

namespace sb2 {

struct ExecScratch {
    double vSafe, vSafeMin, vSafeMinDist;
    bool haveToWait;
};

// MSVehicle::processLinkApproaches; returns false if the right of way at some link depends on the new speed of a foe that is not
// known yet (live == LIVE_PASS1 / LIVE_PASS2, see sb2_link.h): the caller defers the vehicle, nothing has been written
SB2_HD bool veh_processLinkApproaches(const Ctx& c, int v, int wp, const CF& cf, double speed, double pos, int myLane, SUMOTime waiting,
                                      double& vSafe, double& vSafeMin, double& vSafeMinDist, bool& haveToWait,
                                      SUMOTime& jentry, SUMOTime& jconflict, int live) {
    const int cur = c.cur;
    haveToWait = false;
    bool canBrakeVSafeMin = false;
    const int n = c.veh_ndi[cur][wp];
    const double vehLength = cf.length();
    for (int i = 0; i < n; i++) {
        const size_t r = (size_t)i * c.nV + wp;
        const int link = c.di_link[cur][r];
        const bool setRequest = c.di_request[cur][r] != 0;
        const double dpiDist = c.di_dist[cur][r];
        const double vLinkWait = c.di_vwait[cur][r];
        const double vLinkPass = c.di_vpass[cur][r];
        if (link >= 0 && setRequest) {
            const uint8_t ls = c.link_state[link];
            const bool yellow = state_yellow(ls);
            const bool canBrake = (dpiDist > cf_brakeGap(cf, speed, cf.decel(), 0.)
                                   || (speed < cf.decel() * c.TS));
            const bool beyondStopLine = dpiDist < 0.;   // getVehicleStopOffset == 0
            const bool ignoreRedLink = false || beyondStopLine;
            if (yellow && canBrake && !ignoreRedLink) {
                vSafe = vLinkWait;
                haveToWait = true;
                break;
            }
            const double accelV = c.di_accelv[cur][r];
            const double leaveSpeed = accelV < 0 ? vLinkPass : accelV;
            bool opened = yellow;
            if (!opened) {
                const int o = link_opened(c, link, cur, c.di_arrtime[cur][r], c.di_arrspeed[cur][r], leaveSpeed,
                                          vehLength, canBrake ? veh_impatience(c, v) : 1, cf.decel(), waiting, ignoreRedLink, v, live);
                if (o == OPEN_UNDETERMINED) {
                    return false;
                }
                opened = o == OPEN_YES;
            }
            const double visibilityDistance = c.link_foevis[link];
            const bool determinedFoePresence = dpiDist <= visibilityDistance;
            if (opened && !state_priority(ls) && !link_lastWasContMajor(c, link) && !link_isCont(c, link) && !ignoreRedLink) {
                if (!determinedFoePresence && (canBrake || !yellow)) {
                    vSafe = vLinkWait;
                    haveToWait = true;
                    break;
                } else {
                    vSafeMinDist = dpiDist;
                    vSafeMin = MIN3((double)((vSafeMinDist + SB2_POSITION_EPS) / c.TS), vLinkPass, cf_maxNextSpeed(cf, speed));
                    canBrakeVSafeMin = canBrake;
                }
            }
            if (opened) {
                vSafe = vLinkPass;
                if (vSafe < cf.decel() && vSafe <= vLinkWait && vSafe < cf_maxNextSpeed(cf, speed)) {
                    haveToWait = true;
                }
            } else if (ls == 'Z') {
                raise(c, ERR_UNSUPPORTED_LINKSTATE);
            } else if (!canBrake
                       && c.link_tl[link] < 0
                       && dpiDist < cf_brakeGap(cf, speed, cf.emergencyDecel(), 0.)) {
                vSafe = vLinkPass;
            } else {
                vSafe = vLinkWait;
                haveToWait = true;
                break;
            }
            if (lane_internal(c, myLane) && jentry == SB2_TIME_MAX) {
                jentry = c.t;
            }
        } else {
            if (link >= 0 && c.link_internal_before[link] >= 0 && lane_internal(c, myLane) && c.link_junction[link] == c.lane_to_junction[myLane]) {
                jentry = SB2_TIME_MAX;
                jconflict = SB2_TIME_MAX;
            }
            vSafe = vLinkWait;
            // myStopDist == max (no stops): the condition reduces to link != nullptr
            if (link >= 0) {
                if (vSafe < speed) {
                    haveToWait = true;
                } else if (vSafe < SB2_HALTING_SPEED) {
                    haveToWait = true;
                }
            }
            break;
        }
    }
    if (vSafe + SB2_NUMERICAL_EPS < vSafeMin) {
        if (canBrakeVSafeMin && vSafe < speed) {
            vSafe = MIN2(vSafe, MAX2(cf_minNextSpeed(cf, speed), cf_stopSpeed(cf, speed, vSafeMinDist)));
            vSafeMin = 0;
            haveToWait = true;
        } else {
            vSafeMin = vSafe;
        }
    }
    if (c.lane_flags[myLane] & LF_ROUNDABOUT) {
        haveToWait = false;
    }
    (void)pos;
    return true;
}

// MSCFModel::finalizeSpeed (MSCFModel.cpp:199-265) with Krauss patchSpeedBeforeLC; `random` is the lane-RNG draw
SB2_HD double cf_finalizeSpeed(const Ctx& c, const CF& cf, int v, int lane, double oldV, double vPos, SUMOTime timeSinceStartup, double random) {
    const double vStop = MIN2(vPos, vPos);
    const double vMinEmergency = cf_minNextSpeedEmergency(cf, oldV);
    const double vMin = MIN2(cf_minNextSpeed(cf, oldV), MAX2(vPos, vMinEmergency));
    const double factor = 1.;
    double aMax = (MAX2(lane_vehicle_max_speed(c, lane, v), vPos) * factor - oldV) / c.TS;
    double vMax = MIN3(oldV + aMax * c.TS, cf_maxNextSpeed(cf, oldV), vStop);
    vMax = MAX2(vMin, vMax);
    double vNext = MAX2(vMin, cf_dawdle2(cf, vMax, cf.sigma(), random));
    // MSLCM_LC2013::patchSpeed is the identity while the vehicle carries no lane-change state (single-lane edges)
    // applyStartupDelay (MSCFModel.cpp:269-285)
    if (timeSinceStartup > 0 && timeSinceStartup - c.DELTA_T < cf.startupDelay()) {
        const SUMOTime remainingDelay = cf.startupDelay() - (timeSinceStartup - c.DELTA_T);
        if (remainingDelay >= c.DELTA_T) {
            vNext = 0;
        } else {
            vNext = (double)(c.DELTA_T - remainingDelay) / (double)c.DELTA_T * vNext;
        }
    }
    return vNext;
}

// the [0,1) draw of the dawdle step.  The reference draws from one of 64 per-lane-bucket mt19937 streams in active-lane order
// (RandHelper.cpp:102-106, MSLane.cpp:296); the engine uses a counter-based generator (splitmix64 of seed, vehicle, step), so
// sigma > 0 runs are statistically, not bitwise, equivalent unless the draws are replayed (rng_mode 2).
SB2_HD double veh_random(const Ctx& c, int v) {
    if (c.rng_mode == 0) {
        return 0.;
    }
    const int ext = c.veh_ext[v];   // keyed by the vehicle's external index: independent of where its state currently lies
    if (c.rng_mode == 2) {
        return c.rng_replay[ext];
    }
    uint64_t z = c.rng_seed + 0x9E3779B97F4A7C15ull * ((uint64_t)ext + 1) + 0xD1B54A32D192ED03ull * (uint64_t)(c.t / c.DELTA_T + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (double)(z >> 32) / 4294967296.0;   // 32 random bits like rng() / 4294967296.0
}

// MSVehicle::executeMove; `random` = RandHelper::rand(getRNG()) for the dawdle step.  Returns false if the vehicle had to be
// deferred to the second pass (nothing written; see LIVE_PASS1 in sb2_link.h)
SB2_HDN bool veh_executeMove(const Ctx& c, int v, int wp, double random, int live = LIVE_PASS1) {
    int myLane = c.veh_lane[v];
    c.veh_moved[v] = 0;
    if (myLane < 0) {
        return true;
    }
    const int cur = c.cur;
    const int vt = c.veh_type[v];
    CF cf = make_cf(c, vt);
    cf.desSpeed = lane_vehicle_max_speed(c, myLane, v);
    const double oldSpeed = c.veh_speed[v];
    double pos = c.veh_pos[v];
    SUMOTime waiting = c.veh_waiting[v];
    SUMOTime jentry = c.veh_jentry[v], jentryNY = c.veh_jentry_ny[v], jconflict = c.veh_jconflict[v];

    double vSafe = SB2_DBL_MAX;
    double vSafeMin = -SB2_DBL_MAX;
    double vSafeMinDist = 0;
    bool haveToWait = false;
    if (!veh_processLinkApproaches(c, v, wp, cf, oldSpeed, pos, myLane, waiting, vSafe, vSafeMin, vSafeMinDist, haveToWait, jentry, jconflict, live)) {
        return false;
    }

    double vNext = vSafe;
    const double rawAccel = (MAX2(vNext, 0.) - oldSpeed) / c.TS;
    SUMOTime timeSinceStartup = c.veh_since_startup[v];
    if (vNext <= SB2_HALTING_SPEED * c.TS && waiting > c.startupWaitThreshold && rawAccel <= 0.5 * cf.accel()) {
        timeSinceStartup = 0;
    } else {
        timeSinceStartup += c.DELTA_T;
    }
    vNext = cf_finalizeSpeed(c, cf, v, myLane, oldSpeed, vSafe, timeSinceStartup, random);
    if (vNext > 0) {
        vNext = MAX2(vNext, vSafeMin);
    }
    // myStopDist == DBL_MAX > POSITION_EPS
    if (fabs(vNext) < SB2_NUMERICAL_EPS_SPEED) {
        vNext = 0.;
    }
    vNext = MAX2(vNext, 0.);
    if (c.traci_speed_used && c.veh_traci_speed[v] >= 0.) {
        // MSVehicle::processTraCISpeedControl (MSVehicle.cpp:4025-4056)
        vNext = traci_influenceSpeed(c.veh_traci_speed[v], vSafe, MAX2(0., cf_minNextSpeed(cf, oldSpeed)), cf_maxNextSpeed(cf, oldSpeed));
    }

    // updateState (MSVehicle.cpp:4965-5018)
    const double deltaPos = vNext * c.TS;
    const double accel = (MAX2(vNext, 0.) - oldSpeed) / c.TS;
    c.veh_prevspeed[v] = oldSpeed;
    double speed = MAX2(vNext, 0.);
    pos += deltaPos;
    double lastDist = deltaPos;
    // updateWaitingTime (MSVehicle.cpp:4291-4302)
    if (vNext <= SB2_HALTING_SPEED && accel <= 0.5 * cf.accel()) {
        waiting += c.DELTA_T;
    } else {
        waiting = 0;
    }

    // processLaneAdvances (MSVehicle.cpp:4469-4628)
    int32_t* further = c.veh_further + (size_t)v * MAX_FURTHER;
    const int nfOld = c.veh_nfurther[v];
    int passed[MAX_FURTHER + MAX_DRIVE_ITEMS + 2];
    int np = 0;
    for (int i = nfOld - 1; i >= 0; --i) {
        passed[np++] = further[i];
    }
    if (np == 0 || passed[np - 1] != myLane) {
        passed[np++] = myLane;
    }
    int routePos = c.veh_routepos[v];
    const int routeLen = veh_route_len(c, v);
    const int oldLane = myLane;
    double odometer = c.veh_odometer[v];
    double tripRouteLen = c.veh_trip_routelen[v];
    bool arrived = false;
    bool emergency = false;
    int arrivalLaneEntered = -1;
    if (pos > c.lane_length[myLane]) {
        if (routePos != routeLen - 1) {
            int approachedLane = myLane;
            int item = 0;
            const int n = c.veh_ndi[cur][wp];
            while (item < n && approachedLane >= 0 && pos > c.lane_length[approachedLane]) {
                const size_t r = (size_t)item * c.nV + wp;
                const int link = c.di_link[cur][r];
                ++item;
                if (link >= 0) {
                    approachedLane = c.link_via[link] >= 0 ? c.link_via[link] : c.link_to[link];
                    if (state_red(c.link_state[link]) /* && !ignoreRed && !beyondStopLine(dist<0) */ && !(c.di_dist[cur][r] < 0.)) {
                        emergency = true;   // " because of a red traffic light"
                        break;
                    }
                } else if (pos < c.lane_length[myLane] + SB2_NUMERICAL_EPS) {
                    approachedLane = myLane;
                    pos = c.lane_length[myLane];
                } else {
                    raise(c, ERR_NO_CONNECTION);
                    approachedLane = -1;
                    break;
                }
                if (approachedLane != myLane && approachedLane >= 0) {
                    // leaveLane(NOTIFICATION_JUNCTION); MSDevice_Tripinfo::notifyLeave adds the length of the lane that is left
                    tripRouteLen += c.lane_length[myLane];
                    odometer += c.lane_length[myLane];
                    pos -= c.lane_length[myLane];
                    // enterLaneAtMove
                    veh_e1_enter_at_move(c, v, myLane, approachedLane);
                    veh_e2_enter_at_move(c, v, myLane, approachedLane);
                    md_count(c, myLane, MD_LEFT);               // leaveLane(NOTIFICATION_JUNCTION) -> notifyLeave of the lane that is left
                    md_count(c, approachedLane, MD_ENTERED);    // activateReminders(NOTIFICATION_JUNCTION) -> notifyEnter of the entered lane
                    myLane = approachedLane;
                    if (!lane_internal(c, myLane)) {
                        ++routePos;
                    }
                    const int lf = c.link_flags[link];
                    if (lf & KF_ENTRY) {
                        jentry = c.t;
                        jentryNY = jentry;
                    }
                    if (lf & KF_CONFLICT_ENTRY) {
                        jconflict = c.t;
                        jentry = jentryNY;
                    }
                    if (lf & KF_EXIT) {
                        jentry = SB2_TIME_MAX;
                        jentryNY = SB2_TIME_MAX;
                        jconflict = SB2_TIME_MAX;
                    }
                    // hasArrivedInternal
                    if (routePos == routeLen - 1 && pos > MIN2(c.lane_length[myLane], c.veh_arrivalpos[v]) - SB2_POSITION_EPS) {
                        arrivalLaneEntered = approachedLane;   // (not a "passed" lane, but its move reminders were activated)
                        break;
                    }
                    if (np < MAX_FURTHER + MAX_DRIVE_ITEMS + 2) {
                        passed[np++] = approachedLane;
                    }
                }
            }
        } else if (!(pos > MIN2(c.lane_length[myLane], c.veh_arrivalpos[v]) - SB2_POSITION_EPS) && pos < c.lane_length[myLane] + SB2_NUMERICAL_EPS) {
            pos = c.lane_length[myLane];
        }
    }
    arrived = (routePos == routeLen - 1 && pos > MIN2(c.lane_length[myLane], c.veh_arrivalpos[v]) - SB2_POSITION_EPS);

    // updateTimeLoss (MSVehicle.cpp:4306-4315)
    {
        const double vmax = MIN2(lane_vehicle_max_speed(c, myLane, v), MAX2(c.veh_stopspeed[v], vNext));
        if (vmax > 0) {
            c.veh_timeloss[v] += c.TS * (vmax - vNext) / vmax;
        }
    }
    double accelOut = accel;
    int nfNew = 0;
    double backPos = c.veh_backpos[v];
    if (!arrived) {
        if (pos > c.lane_length[myLane]) {
            // emergency stop at the end of the lane (MSVehicle.cpp:4784-4796)
            (void)emergency;
            raise(c, ERR_EMERGENCY_STOP);
            count_event(c, oldLane, 5);
            pos = c.lane_length[myLane];
            speed = 0;
            accelOut = 0;
        }
        // updateFurtherLanes (MSVehicle.cpp:5055-5131)
        double backPosOnPreviousLane = pos - cf.length();
        int newFurther[MAX_FURTHER];
        if (np > 1) {
            for (int pi = np - 2; pi >= 0 && backPosOnPreviousLane < 0; --pi) {
                if (nfNew < MAX_FURTHER) {
                    newFurther[nfNew++] = passed[pi];
                } else {
                    raise(c, ERR_FURTHER_OVERFLOW);
                }
                backPosOnPreviousLane += c.lane_length[passed[pi]];
            }
        }
        for (int i = 0; i < nfNew; i++) {
            further[i] = newFurther[i];
        }
        backPos = backPosOnPreviousLane;
    }
    // ---- write back
    c.veh_pos[v] = pos;
    c.veh_speed[v] = speed;
    c.veh_accel[v] = accelOut;
    c.veh_lastdist[v] = lastDist;
    c.veh_backpos[v] = backPos;
    c.veh_nfurther[v] = nfNew;
    c.veh_waiting[v] = waiting;
    c.veh_since_startup[v] = timeSinceStartup;
    c.veh_jentry[v] = jentry;
    c.veh_jentry_ny[v] = jentryNY;
    c.veh_jconflict[v] = jconflict;
    c.veh_havetowait[v] = haveToWait;
    c.veh_routepos[v] = routePos;
    c.veh_odometer[v] = odometer;
    c.veh_trip_routelen[v] = tripRouteLen;
    // MSDevice_Tripinfo::notifyMove (not stopped): waiting = slow and not accelerating (accelThresholdForWaiting, MSVehicle.h:2100)
    if (speed <= SB2_HALTING_SPEED && accelOut <= 0.5 * cf.accel()) {
        c.veh_trip_wait[v] += c.DELTA_T;
        if (!c.veh_trip_amwaiting[v]) {
            c.veh_trip_waitcount[v] += 1;
            c.veh_trip_amwaiting[v] = 1;
        }
    } else {
        c.veh_trip_amwaiting[v] = 0;
    }
    veh_emissions_notifyMove(c, v, speed, accelOut);   // workOnMoveReminders (MSVehicle.cpp:4902) -> MSDevice_Emissions::notifyMove
    veh_e1_move(c, v, pos - lastDist, pos, speed);     //                                          -> MSInductLoop::notifyMove
    if (SB2_OBSERVERS && c.nE2 > 0) {
        veh_e2_move(c, v, pos - lastDist, pos, speed, lane_vehicle_max_speed(c, myLane, v));   //       -> MSE2Collector::notifyMove
    }
    if (SB2_OBSERVERS && c.md_on) {
        // -> MeanDataValues::notifyMove of every lane the vehicle's body was on or entered; rem->second = lengths of the lanes left since
        const double vmaxNow = lane_vehicle_max_speed(c, myLane, v);
        const int nChain = np + (arrivalLaneEntered >= 0 ? 1 : 0);
        for (int k = 0; k < nChain; k++) {
            double off = 0.;
            for (int j = k; j < nChain - 1; j++) {
                off += c.lane_length[passed[j]];
            }
            const int lane = k < np ? passed[k] : arrivalLaneEntered;
            md_notifyMove(c, v, lane, (pos - lastDist) + off, pos + off, MAX2(0., speed), vmaxNow);
        }
        if (arrived) {
            md_count(c, myLane, MD_ARRIVED);
        }
    }
    if (arrived) {
        veh_e1_leave_net(c, v);        // onRemovalFromNet(NOTIFICATION_ARRIVED) -> leaveLane -> notifyLeave
        veh_e2_leave_net(c, v);
        c.veh_moved[v] = 2;
        c.veh_arrival_time[v] = c.t;   // MSDevice_Tripinfo::notifyLeave(NOTIFICATION_ARRIVED): getCurrentTimeStep()
        c.veh_arrival_lane[v] = myLane;
        c.lane_arrived[oldLane] = 1;   // its queue entry is removed by lane_handoff_remove wherever it sits; veh_lane becomes -2 there
        count_event(c, oldLane, 1);
    } else if (myLane != oldLane) {
        c.veh_moved[v] = 1;
        c.veh_lane[v] = myLane;
        c.veh_next_incoming[v] = atomic_exch(&c.lane_incoming[myLane], v);
    }
    return true;
}

// first pass of executeMove for the vehicle at work-list position wp; a deferred vehicle is queued for veh_execute_deferred
SB2_HDN void veh_executeMove_pass1(const Ctx& c, int wp) {
    const int v = c.veh_order[c.cur][wp];
    if (veh_executeMove(c, v, wp, veh_random(c, v), LIVE_PASS1)) {
        veh_register_partials(c, v);
    } else {
        c.veh_deferred[v] = 1;
#if defined(__CUDA_ARCH__)
        const int at = atomicAdd(&c.counters[9], 1);
#else
        const int at = c.counters[9]++;
#endif
        c.deferred_list[at] = wp;
    }
}

// one round of the second pass for entry i of the deferred list; returns true if the vehicle is (now) done
SB2_HDN bool veh_execute_deferred(const Ctx& c, int i) {
    const int wp = c.deferred_list[i];
    const int v = c.veh_order[c.cur][wp];
    if (!c.veh_deferred[v]) {
        return true;
    }
    if (!veh_executeMove(c, v, wp, veh_random(c, v), LIVE_PASS2)) {
        return false;
    }
    veh_register_partials(c, v);
#if defined(__CUDA_ARCH__)
    __threadfence();
#endif
    c.veh_deferred[v] = 0;
    return true;
}

}  // namespace sb2
