/* sb2_plan.h -- MSVehicle::planMove for one vehicle (one GPU thread per vehicle).
 *
 * Restates src/microsim/MSVehicle.cpp:2162-2229 (planMove), 2276-3122 (planMoveInternal; branches for stops, rail,
 * opposite driving, sublanes, pedestrians, influencers are out of scope and refused at load), 3169-3184 (getArrivalTime),
 * 3188-3265 (adaptToLeaders), 3300-3382 (adaptToLeader), 3386-3496 (adaptToJunctionLeader), 3568-3718 (checkLinkLeader),
 * 5255-5489 (checkRewindLinkLanes), 5493-5506 (setApproachingForAllLinks), 7681-7821 (isLeader);
 * src/microsim/MSLink.cpp:1583-1907 (getLeaderInfo, vehicle foes only); src/microsim/MSLane.cpp:1588-1624 (planMovements).
 *
 * Everything read from other vehicles is state of the previous step (positions, speeds, junction entry times,
 * myHaveToWaitOnNextLink, the previous approach table), so all vehicles plan in parallel.
 */
#pragma once
#include "sb2_types.h"
#include "sb2_cf.h"
#include "sb2_veh.h"
#include "sb2_link.h"

namespace sb2 {

struct DriveItem {
    int link;
    double vPass, vWait;
    bool setRequest;
    SUMOTime arrivalTime;
    double arrivalSpeed, arrivalSpeedBraking, distance, accelV;
    bool hadStoppedVehicle;
    double availableSpace;
};
SB2_HD void di_adaptLeaveSpeed(DriveItem& d, double v) { d.accelV = d.accelV < 0 ? v : MIN2(d.accelV, v); }
SB2_HD double di_getLeaveSpeed(const DriveItem& d) { return d.accelV < 0 ? d.vPass : d.accelV; }
SB2_HD DriveItem di_make(int link, double vPass, double vWait, bool setRequest, SUMOTime arrivalTime, double arrivalSpeed,
                         double arrivalSpeedBraking, double distance, double leaveSpeed) {
    DriveItem d;
    d.link = link; d.vPass = vPass; d.vWait = vWait; d.setRequest = setRequest; d.arrivalTime = arrivalTime;
    d.arrivalSpeed = arrivalSpeed; d.arrivalSpeedBraking = arrivalSpeedBraking; d.distance = distance; d.accelV = leaveSpeed;
    d.hadStoppedVehicle = false; d.availableSpace = 0;
    return d;
}
// DriveProcessItem(double vWait, double distance, double availableSpace = 0), MSVehicle.h:2014
SB2_HD DriveItem di_make_nolink(double vWait, double distance, double availableSpace) {
    DriveItem d = di_make(-1, vWait, vWait, false, 0, 0, 0, distance, -1);
    d.availableSpace = availableSpace;
    return d;
}

struct Ego {
    int v, vt, lane;
    double pos, speed, length, minGap;
    CF cf;
    SUMOTime waiting;
};

// MSVehicle::isLeader (MSVehicle.cpp:7681-7821)
SB2_HD bool veh_isLeader(const Ctx& c, const Ego& e, int link, int veh, double gap) {
    if (veh < 0) {
        return false;
    }
    if (!lane_internal(c, e.lane) || c.lane_to_junction[e.lane] != c.link_junction[link]) {
        return true;
    }
    const int foeLane = c.veh_lane[veh];
    if (lane_internal(c, foeLane)) {
        if (c.lane_from_junction[foeLane] == c.link_junction[link]) {
            SUMOTime egoET = c.veh_jconflict[e.v];
            SUMOTime foeET = c.veh_jentry[veh];
            if (c.lane_pred_normal[foeLane] == c.lane_pred_normal[c.link_internal_before[link]]) {
                egoET = c.veh_jentry_ny[e.v];
                foeET = c.veh_jentry_ny[veh];
                if ((c.link_flags[link] & KF_EXIT_AFTER_IJ)
                        && (c.link_flags[c.lane_entry_link[c.lane_pred_logical[c.link_internal_before[link]]]] & KF_INDIRECT)) {
                    egoET = c.veh_jconflict[e.v];
                }
            } else {
                const int foeLink = c.inc_link[c.lane_inc_begin[foeLane]];
                bool response, response2;
                const int entry = c.link_entry[link];
                const int foeEntry = c.link_entry[foeLink];
                const uint8_t es = c.link_state[entry], fs = c.link_state[foeEntry];
                if (state_red(es) || state_red(fs)) {
                    if (!state_red(fs) && c.veh_speed[veh] > SB2_HALTING_SPEED && gap < 0) {
                        const int ft = c.veh_type[veh];
                        const CF fcf = make_cf(c, ft);
                        const double foeNextSpeed = c.veh_speed[veh] + fcf.accel() * c.TS;
                        const double foeBrakeGap = cf_brakeGap(fcf, foeNextSpeed, fcf.decel(), fcf.tau());
                        const double foeGap = -gap - c.vt_length[ft] - 2 * e.minGap;
                        if (foeGap < foeBrakeGap) {
                            response = true; response2 = false;
                        } else {
                            response = false; response2 = true;
                        }
                    } else {
                        response = true; response2 = true;
                    }
                } else if (state_priority(es) != state_priority(fs)) {
                    response = !state_priority(es);
                    response2 = !state_priority(fs);
                } else if (state_yellow(es) && state_yellow(fs)) {
                    response = c.veh_speed[veh] >= e.speed;
                    response2 = e.speed >= c.veh_speed[veh];
                } else {
                    const int li = c.link_index[link], fi = c.link_index[foeLink];
                    response = (c.link_response[(size_t)link * 4 + (fi >> 6)] >> (fi & 63)) & 1;
                    response2 = (c.link_response[(size_t)foeLink * 4 + (li >> 6)] >> (li & 63)) & 1;
                }
                if (!response) {
                    foeET = c.veh_jconflict[veh];
                    egoET = c.veh_jentry[e.v];
                } else if (response && response2) {
                    foeET = c.veh_jconflict[veh];
                    egoET = c.veh_jconflict[e.v];
                }
            }
            if (egoET == foeET) {
                if (e.speed == c.veh_speed[veh]) {
                    return c.veh_idrank[e.v] < c.veh_idrank[veh];   // getID() < veh->getID()
                } else {
                    return e.speed < c.veh_speed[veh];
                }
            } else {
                return egoET > foeET;
            }
        } else {
            return true;
        }
    } else {
        return true;
    }
}

// MSVehicle::adaptToLeader (MSVehicle.cpp:3300-3382); lastLink may be null
SB2_HD void veh_adaptToLeader(const Ctx& c, const Ego& e, int pred, double gap, double seen, DriveItem* lastLink, double& v, double& vLinkPass) {
    if (pred < 0) {
        return;
    }
    double vsafeLeader = 0;
    bool backOnRoute = true;
    if (gap < 0 && lastLink != nullptr && lastLink->link >= 0) {
        backOnRoute = false;
        const int lk = lastLink->link;
        const int current = c.link_via[lk] >= 0 ? c.link_via[lk] : c.link_to[lk];
        const int predBackLane = veh_back_lane(c, pred);
        if (predBackLane == current) {
            backOnRoute = true;
        } else {
            // walk getBestLanesContinuation(): the normal lanes along the route (MSVehicle.cpp:6208-6214 limits it to
            // about 8 edges), stopping at `current`
            int lane = e.lane;
            int view = 0;
            if (lane_internal(c, lane)) {
                lane = -1;   // internal lanes are not kept in the bestLanes structure (placeholder entry)
            }
            for (int guard = 0; guard < 9; guard++) {
                if (lane >= 0) {
                    if (lane == current) {
                        break;
                    }
                    if (predBackLane == lane) {
                        backOnRoute = true;
                    }
                }
                const int from = lane >= 0 ? lane : e.lane;
                const int k = lane_succ_link(c, from, veh_succ_edge(c, e.v, view + 1));
                if (k < 0) {
                    break;
                }
                lane = c.link_to[k];
                view++;
            }
        }
        if (!backOnRoute) {
            double stopDist = seen - c.lane_length[current] - SB2_POSITION_EPS;
            if (c.link_internal_before[lk] >= 0) {
                stopDist -= c.lane_length[c.link_internal_before[lk]];
            }
            vsafeLeader = cf_stopSpeed(e.cf, e.speed, stopDist);
        }
    }
    const double predSpeed = c.veh_speed[pred];
    const double predDecel = c.vt_apparentdecel[c.veh_type[pred]];
    if (backOnRoute) {
        vsafeLeader = cf_followSpeed(e.cf, e.speed, gap, predSpeed, predDecel);
    }
    if (lastLink != nullptr) {
        const double futureVSafe = cf_followSpeed(e.cf, lastLink->accelV, gap, predSpeed, predDecel);
        di_adaptLeaveSpeed(*lastLink, futureVSafe);
    }
    v = MIN2(v, vsafeLeader);
    vLinkPass = MIN2(vLinkPass, vsafeLeader);
}

// MSVehicle::adaptToLeaders for one sublane (MSVehicle.cpp:3188-3265)
SB2_HD void veh_adaptToLeaders(const Ctx& c, const Ego& e, int pred, double seen, DriveItem* lastLink, int lane, double& v, double& vLinkPass) {
    if (pred >= 0 && pred != e.v) {
        const double predBack = veh_back_pos_on_lane(c, pred, lane);
        const double gap = (lastLink == nullptr
                            ? predBack - e.pos - e.minGap
                            : predBack + seen - c.lane_length[lane] - e.minGap);
        veh_adaptToLeader(c, e, pred, gap, seen, lastLink, v, vLinkPass);
    }
}

// MSVehicle::adaptToJunctionLeader (MSVehicle.cpp:3386-3496), hasDeparted() == true
SB2_HD void veh_adaptToJunctionLeader(const Ctx& c, const Ego& e, int leader, double gap, double seen, DriveItem* lastLink, int lane,
                                      double& v, double& vLinkPass, double distToCrossing, bool departed = true) {
    if (leader < 0) {
        return;
    }
    double vsafeLeader = 0;
    const double lSpeed = c.veh_speed[leader];
    if (gap >= 0) {
        if (departed) {
            vsafeLeader = cf_followSpeed(e.cf, e.speed, gap, lSpeed, c.vt_apparentdecel[c.veh_type[leader]]);
        } else {
            vsafeLeader = cf_insertionFollowSpeed(e.cf, e.speed, gap, lSpeed, c.vt_apparentdecel[c.veh_type[leader]]);
        }
    } else if (leader != e.v) {
        vsafeLeader = cf_stopSpeed(e.cf, e.speed, seen - c.lane_length[lane] - SB2_POSITION_EPS);
    }
    if (distToCrossing >= 0) {
        const double vStop = cf_stopSpeed(e.cf, e.speed, distToCrossing - e.minGap);
        if (leader == e.v) {
            // pedestrians: out of scope
        } else if (gap == -SB2_DBL_MAX) {
            vsafeLeader = MAX2(vsafeLeader, vStop);
        } else {
            const double leaderDistToCrossing = distToCrossing - gap;
            const double leaderPastCPTime = leaderDistToCrossing / MAX2(lSpeed, SB2_HALTING_SPEED);
            const double vFinal = MAX2(e.speed, 2 * (distToCrossing - e.minGap) / leaderPastCPTime - e.speed);
            const double v2 = e.speed + ((vFinal - e.speed) / leaderPastCPTime) * c.TS;
            vsafeLeader = MAX2(vsafeLeader, MIN2(v2, vStop));
        }
    }
    if (lastLink != nullptr) {
        di_adaptLeaveSpeed(*lastLink, vsafeLeader);
    }
    v = MIN2(v, vsafeLeader);
    vLinkPass = MIN2(vLinkPass, vsafeLeader);
}

// MSVehicle::checkLinkLeader fused with MSLink::getLeaderInfo (each link leader is consumed as soon as it is found; the
// reference builds the list first, MSVehicle.cpp:3592, but nothing it reads changes in between).
// `egoOnRoad` false = called from MSLane::isInsertionSuccess (MSLane.cpp:1111).
SB2_HD void veh_checkLinkLeader(const Ctx& c, const Ego& e, int link, int lane, double seen, DriveItem* lastLink,
                                double& v, double& vLinkPass, double& vLinkWait, bool& setRequest, int tableBuf, bool departed = true) {
    const int ilb = c.link_internal_before[link];
    if (ilb >= 0) {   // fromInternalLane()
        const int myLaneTo = c.link_to[link];
        const bool isIJ = (c.link_flags[link] & KF_INTERNAL_JUNCTION) != 0;
        const bool isExitAfterIJ = (c.link_flags[link] & KF_EXIT_AFTER_IJ) != 0;
        const double radius = c.link_radius[link];
        for (int i = c.link_foelane_begin[link]; i < c.link_foelane_begin[link + 1]; ++i) {
            const int foeLane = c.foelane[i];
            if (foeLane < 0) {
                continue;   // a conflict record without foe lane (only addressed through conf_foeidx, MSLink.cpp:614)
            }
            if (c.lane_last[foeLane] < 0 && c.lane_npartial[foeLane] == 0) {
                continue;   // nobody on the foe lane: the AnyVehicleIterator loop below would not run (nothing above it has side effects)
            }
            const int foeExitLink = c.lane_link_begin[foeLane];
            const int cflag = c.conf_flag[i];
            // ConflictInfo::getLengthBehindCrossing / getFoeLengthBehindCrossing / getFoeConflictSize (MSLink.cpp:83-109)
            const double lbc = (cflag == CONFLICT_STOP_AT_INTERNAL_JUNCTION) ? c.lane_length[ilb] : c.conf_lbc[i];
            const int fci = c.conf_foeidx[i];
            const int foeConfBase = c.link_foelane_begin[foeExitLink];
            const double foeLbc = (cflag == CONFLICT_DUMMY_MERGE) ? 0 : (fci >= 0 ? c.conf_lbc[foeConfBase + fci] : -SB2_NO_INTERSECTION);
            double distToCrossing = seen - lbc;
            const double foeDistToCrossing = c.lane_length[foeLane] - foeLbc;
            const bool foeExitIsIJ = (c.link_flags[foeExitLink] & KF_INTERNAL_JUNCTION) != 0;
            const bool sameTarget = (myLaneTo == c.link_to[foeExitLink]) && !isIJ && !foeExitIsIJ;
            const bool sameSource = (c.lane_pred_normal[ilb] == c.lane_pred_normal[foeLane]);
            const double crossingWidth = (sameTarget || sameSource) ? 0 : c.conf_size[i];
            const double foeCrossingWidth = (sameTarget || sameSource) ? 0 : (fci >= 0 ? c.conf_size[foeConfBase + fci] : 0);
            const int foeExitNext = c.link_via[foeExitLink] >= 0 ? c.link_via[foeExitLink] : c.link_to[foeExitLink];
            const bool contLane = (lane_internal(c, foeExitNext) && !(isIJ || isExitAfterIJ));
            if (distToCrossing + crossingWidth < 0 && !sameTarget) {
                continue;
            }
            bool ignoreGreenCont = false;
            bool foeIndirect = false;
            if (contLane) {
                const int entry = c.lane_entry_link[ilb];
                const int foeEntry = c.lane_entry_link[foeLane];
                foeIndirect = (c.link_flags[foeEntry] & KF_INDIRECT) != 0;
                if (entry >= 0 && state_green(c.link_state[entry])
                        && foeEntry >= 0 && state_green(c.link_state[foeEntry])
                        && c.link_from[entry] != c.link_from[foeEntry]) {
                    ignoreGreenCont = true;
                }
            }
            if (foeIndirect && distToCrossing >= SB2_NO_INTERSECTION) {
                continue;
            }
            AnyVehIter it = any_begin(c, foeLane);
            for (int leader = any_next(c, foeLane, it); leader >= 0; leader = any_next(c, foeLane, it)) {
                (void)SB2_CHK(leader, 0, c.nV, 10);
                const int lt = c.veh_type[leader];
                const double leaderLen = c.vt_length[lt];
                const double leaderBack = veh_back_pos_on_lane(c, leader, foeLane) - 0 /*extraGap*/;
                const double leaderBackDist = foeDistToCrossing - leaderBack;
                const double l2 = e.length + 2;
                const double sagitta = radius != SB2_DBL_MAX ? radius - sqrt(radius * radius - 0.25 * l2 * l2) : 0;
                const bool pastTheCrossingPoint = leaderBackDist + foeCrossingWidth + sagitta < 0;
                const bool enteredTheCrossingPoint = leaderBackDist < leaderLen;
                const bool cannotIgnore = (contLane || sameTarget || sameSource);
                const bool inTheWay = ((((!pastTheCrossingPoint && distToCrossing > 0) || (sameTarget && distToCrossing > leaderBackDist - leaderLen))
                                        && (enteredTheCrossingPoint || (sameSource && !enteredTheCrossingPoint && foeDistToCrossing < distToCrossing))
                                        && (!(cflag == CONFLICT_DUMMY_MERGE) || sameSource))
                                       || false /* bidi */);
                const int avi = link_getApproaching(c, foeExitLink, tableBuf, leader);
                const bool aviWillPass = avi >= 0 && c.di_request[tableBuf][avi];
                const bool willPass = aviWillPass || (avi < 0 && sameTarget);
                if (leader == e.v) {
                    continue;
                }
                if (!inTheWay && ignoreGreenCont) {
                    continue;
                }
                if (distToCrossing < -SB2_POSITION_EPS && !inTheWay) {
                    continue;
                }
                if (sameSource
                        && c.lane_edge[e.lane] == c.lane_edge[ilb]
                        && leaderBack + leaderLen < e.pos - e.length) {
                    continue;
                }
                const bool leaderFrontOnLane = veh_front_on_lane(c, leader, foeLane);
                if ((!cannotIgnore || sameTarget)
                        && !willPass
                        && (avi < 0 || c.veh_speed[leader] < SB2_HALTING_SPEED)
                        && leaderFrontOnLane
                        && !inTheWay) {
                    continue;
                }
                if (c.veh_waiting[leader] < c.ignoreJunctionBlocker) {
                    double gap;
                    if (contLane && !sameSource) {
                        gap = -SB2_DBL_MAX;
                    } else {
                        if (pastTheCrossingPoint && !sameTarget) {
                            continue;
                        }
                        double leaderBackDist2 = leaderBackDist;
                        if (sameTarget && leaderBackDist2 < 0) {
                            const double mismatch = foeLbc - lbc;
                            if (mismatch > 0) {
                                leaderBackDist2 += mismatch;
                            }
                        }
                        gap = distToCrossing - e.minGap - leaderBackDist2 - foeCrossingWidth;
                    }
                    const bool stopAsap = (leaderFrontOnLane ? cannotIgnore : (sameTarget || sameSource));
                    const double llDistToCrossing = stopAsap ? -1 : distToCrossing;
                    // ---- MSVehicle::checkLinkLeader body for this link leader (MSVehicle.cpp:3627-3703)
                    if (veh_isLeader(c, e, link, leader, gap) || inTheWay) {
                        veh_adaptToJunctionLeader(c, e, leader, gap, seen, lastLink, lane, v, vLinkPass, llDistToCrossing, departed);
                        if (lastLink != nullptr) {
                            v = MAX2(v, lastLink->vWait);
                        }
                        const int leaderLane = c.veh_lane[leader];
                        if (v < SB2_HALTING_SPEED
                                && (c.lane_pred_logical[leaderLane] == c.lane_pred_logical[e.lane]
                                    || c.lane_pred_logical[leaderLane] == e.lane
                                    || c.veh_waiting[leader] > TIME2STEPS(SB2_JUNCTION_BLOCKAGE_TIME))) {
                            setRequest = false;
                            if (lastLink != nullptr && c.lane_pred_logical[leaderLane] == e.lane) {
                                lastLink->setRequest = false;
                            }
                        }
                    }
                }
            }
        }
    }
    vLinkWait = MIN2(vLinkWait, v);
}

// MSVehicle::getArrivalTime (MSVehicle.cpp:3169-3184)
SB2_HD SUMOTime veh_getArrivalTime(const Ctx& c, const Ego& e, SUMOTime t, double seen, double v, double arrivalSpeed) {
    return t - c.DELTA_T + cf_getMinimalArrivalTime(e.cf, seen, v, arrivalSpeed);
}

// MSVehicle::keepClear (MSVehicle.cpp:7592-7600), jmIgnoreKeepClearTime default -1
SB2_HD bool veh_keepClear(const Ctx& c, int link) { return (c.link_flags[link] & KF_HASFOES) && (c.link_flags[link] & KF_KEEPCLEAR); }

}  // namespace sb2
