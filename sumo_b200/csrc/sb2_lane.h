/* sb2_lane.h -- per-lane work (one GPU thread per lane): pre-plan summary, post-execute hand-off, collision check; and
 * the static traffic-light phase clock (one thread per program).
 *
 * Restates src/microsim/MSLane.cpp:1588-1624 (planMovements: cumulatedVehLength), 4734-4754 (getSpaceTillLastStanding),
 * 2311-2401 + 2470-2487 (executeMovements removal bookkeeping, updateLengthSum), 2598-2621 + 3539-3548
 * (integrateNewVehicles, vehicle_position_sorter), 2625-2629 (sortPartialVehicles), 1826-1835 + 1971-2030
 * (detectCollisions / detectCollisionBetween, non-sublane); src/microsim/traffic_lights/MSTrafficLightLogic.cpp:68-92,
 * 402-414 (SwitchCommand::execute, setTrafficLightSignals), MSSimpleTrafficLightLogic.cpp:80-110 (trySwitch),
 * src/microsim/MSLink.cpp:1325-1336 (setTLState).
 */
#pragma once
#include "sb2_types.h"
#include "sb2_cf.h"
#include "sb2_veh.h"
#include "sb2_e1.h"
#include "sb2_e2.h"
#include "sb2_meandata.h"

namespace sb2 {

// MSEdgeControl::myActiveLanes as keys (see Ctx::lane_actkey).  phase 0 = patchActiveLanes (rank: numerical id, the order of
// myChangedStateLanes), 1 = integrateNewVehicles loop (rank: id order), 2 = changeLanes' toAdd
SB2_HD int64_t lane_make_actkey(const Ctx& c, int l, int phase) {
    const int64_t step = c.t / c.DELTA_T + 1;
    const int64_t rank = phase == 1 ? c.lane_idrank[l] : c.lane_numid[l];
    const int64_t key = ((step * 4 + phase) << 32) | rank;
    return (c.lane_flags[l] & LF_LANECHANGER) ? -key : key;   // haveNeighbors: push_front
}
// begin of the step: patchActiveLanes (MSEdgeControl.cpp:118-134), then the erase of empty lanes in planMovements (:150-153)
SB2_HD void lane_patch_active(const Ctx& c, int l) {
    if (c.lane_changed[l]) {
        c.lane_changed[l] = 0;
        if (c.lane_actkey[l] == 0 && c.lane_nveh[l] > 0) {
            c.lane_actkey[l] = lane_make_actkey(c, l, 0);
        }
    }
    if (c.lane_nveh[l] == 0) {
        c.lane_actkey[l] = 0;
    }
}

// before planMove: cumulated lengths in front of every vehicle, space till last standing vehicle
// orderBase = first slot of this lane in the lane-ordered work list veh_order; firstSlot >= 0: the lane's front vehicle goes
// there instead (k_pre groups the front vehicles of a block's lanes), the followers start at orderBase
SB2_HDN void lane_summary(const Ctx& c, int l, int orderBase, int orderStep = 1, int firstSlot = -1) {
    const int vis = lane_vis(c, l);
    if (vis == LV_NONE) {
        return;
    }
    if (vis >= LV_GHOST) {
        lane_patch_active(c, l);   // (a halo lane's key arrives with its state)
    }
    // front -> back (MSLane.cpp:1608-1623)
    double cum = 0.;
    int guard = 0;
    for (int v = c.lane_first[l]; v >= 0; v = c.veh_behind[v]) {
        (void)SB2_CHK(v, 0, c.nV, 11);
        (void)SB2_CHK(orderBase, 0, c.nV, 12);
        (void)SB2_CHK(guard++, 0, c.lane_nveh[l], 13);
        c.veh_len_in_front[v] = cum;
        c.veh_qlane[v] = l;
        if (vis < LV_GHOST) {
            // halo lane: its vehicles are read, not moved
        } else if (firstSlot >= 0) {
            c.veh_order[c.cur][firstSlot] = v;
            firstSlot = -1;
        } else {
            c.veh_order[c.cur][orderBase] = v;
            orderBase += orderStep;
        }
        cum += veh_length_with_gap(c, v);
    }
    // back -> front (MSLane.cpp:4734-4754); `last != ego` never excludes anything for the lanes ahead of ego
    double lengths = 0;
    bool found = false;
    double ret = 0;
    for (int last = c.lane_last[l]; last >= 0; last = c.veh_ahead[last]) {
        if (c.veh_speed[last] < SB2_HALTING_SPEED && !(c.lane_flags[l] & LF_ROUNDABOUT)) {
            found = true;
            const CF cf = make_cf(c, c.veh_type[last]);
            const double lastBrakeGap = cf_brakeGap(cf, c.veh_speed[last]);
            ret = (c.veh_pos[last] - cf.length()) + lastBrakeGap - lengths;
            break;
        }
        lengths += veh_length_with_gap(c, last);
    }
    if (!found) {
        ret = c.lane_length[l] - lengths;
    }
    c.ls_space[l] = ret;
    c.ls_found_stopped[l] = found;
}

// MSVehicleTransfer::add for a teleporting vehicle that its lane's queue has just dropped (MSVehicleTransfer.cpp:54-77):
// onRemovalFromNet(NOTIFICATION_TELEPORT) -> leaveLane (odometer, further lanes, waiting time := 0), then enterLaneAtMove(first lane
// of the next route edge, onTeleporting) puts it at position 0 of that edge, off the road.  Its approach records stay in the
// table but every reader skips vehicles that are not on the road; partial occupations are dropped in lane_handoff_integrate.
SB2_HD void veh_transfer_add(const Ctx& c, int v, int lane) {
    veh_e1_leave_net(c, v);   // leaveLane(NOTIFICATION_TELEPORT / _TELEPORT_ARRIVED) -> notifyLeave of every move reminder
    veh_e2_leave_net(c, v);
    md_count(c, lane, MD_LEFT);
    md_count(c, lane, MD_TELEPORTED);
    if (c.veh_routepos[v] + 1 >= veh_route_len(c, v)) {
        // "teleports beyond arrival edge": removed from the simulation (NOTIFICATION_TELEPORT_ARRIVED)
        c.veh_arrival_time[v] = c.t;
        c.veh_arrival_lane[v] = lane;
        c.veh_trip_amwaiting[v] = 2;   // (free after the removal) marks NOTIFICATION_TELEPORT_ARRIVED for the trip record
        c.veh_moved[v] = 2;
        c.veh_lane[v] = -2;
        count_event(c, lane, 1);
        return;
    }
    c.veh_odometer[v] += c.lane_length[lane];
    c.veh_trip_routelen[v] += c.lane_length[lane];   // MSDevice_Tripinfo::notifyLeave(NOTIFICATION_TELEPORT)
    c.veh_nfurther[v] = 0;
    c.veh_waiting[v] = 0;
    c.veh_routepos[v] += 1;
    c.veh_pos[v] = 0;
    c.veh_lane[v] = -3;
    // enterLaneAtMove(.., onTeleporting): myPos = 0, activateReminders(NOTIFICATION_TELEPORT) (MSVehicle.cpp:5613-5618)
    veh_e1_add_lane(c, v, c.edge_lane_begin[veh_succ_edge(c, v, 0)], true, 0., -veh_length(c, v));
    veh_e2_add_lane(c, v, c.edge_lane_begin[veh_succ_edge(c, v, 0)], true, -veh_length(c, v));
    md_count(c, c.edge_lane_begin[veh_succ_edge(c, v, 0)], MD_ENTERED);
    c.veh_tp_proceed[v] = -1;
    // partitioned run: the transfer entry belongs to the rank that owns the edge the vehicle now (virtually) sits on, because
    // MSVehicleTransfer::checkInsertions tries to re-insert it there in this very step.  That rank sees the teleport itself: the
    // lane the vehicle was taken from is its own or in its ghost column (the next route edge ends at a neighbour junction).
    if (!lane_auth(c, c.edge_lane_begin[veh_succ_edge(c, v, 0)])) {
        return;
    }
#if defined(__CUDA_ARCH__)
    const int at = atomicAdd(&c.counters[10], 1);
#else
    const int at = c.counters[10]++;
#endif
    c.tp_list[at] = v;
}

// after executeMove, pass 1: pop the vehicles that left at the front (reads veh_behind of the leavers, which pass 2
// overwrites when it links them into their new lane -- hence two passes with a grid-wide barrier in between)
SB2_HDN void lane_handoff_remove(const Ctx& c, int l) {
    // ---- MSLane::executeMovements erase + updateLengthSum
    double toRemove = 0;
    int f = c.lane_first[l];
    int removed = 0;
    while (f >= 0 && (c.veh_lane[f] != l || c.veh_moved[f] == 2)) {
        toRemove += veh_length_with_gap(c, f);
        if (c.veh_moved[f] == 2) {
            c.veh_lane[f] = -2;   // arrived (executeMove leaves veh_lane alone so that it stays readable during the whole pass)
        }
        f = c.veh_behind[f];
        removed++;
    }
    if (c.lane_arrived[l]) {
        // a vehicle that reached its arrivalPos behind the front vehicle (explicit arrivalPos inside the lane): MSLane::executeMovements
        // erases it where it is (MSLane.cpp:2334-2340); unlink every arrived entry behind the new front, front -> back like the reference
        c.lane_arrived[l] = 0;
        int ahead = f;
        for (int x = f >= 0 ? c.veh_behind[f] : -1; x >= 0;) {
            const int next = c.veh_behind[x];
            if (c.veh_lane[x] != l || c.veh_moved[x] == 2) {
                if (c.veh_moved[x] == 2) {
                    c.veh_lane[x] = -2;
                }
                toRemove += veh_length_with_gap(c, x);
                removed++;
                c.veh_behind[ahead] = next;
                if (next >= 0) {
                    c.veh_ahead[next] = ahead;
                } else {
                    c.lane_last[l] = ahead;
                }
            } else {
                ahead = x;
            }
            x = next;
        }
    }
    if (removed > 0) {
        c.lane_first[l] = f;
        if (f >= 0) {
            c.veh_ahead[f] = -1;
        } else {
            c.lane_last[l] = -1;
        }
        c.lane_nveh[l] -= removed;
    }
    double brutto = c.lane_brutto[l] - toRemove;
    if (c.lane_nveh[l] == 0) {
        brutto = 0;
    }
    c.lane_brutto[l] = brutto;
    // MSLane.cpp:2402-2456: the front-most remaining vehicle (firstNotStopped; there are no stops on this path) is teleported after
    // waiting longer than time-to-teleport (r1; the highway / disconnected / bidi / rail-signal rules need options that are refused)
    if (f >= 0 && c.timeToTeleport > 0 && c.veh_waiting[f] > c.timeToTeleport) {
        const int next = c.veh_behind[f];
        c.lane_first[l] = next;
        if (next >= 0) {
            c.veh_ahead[next] = -1;
        } else {
            c.lane_last[l] = -1;
        }
        c.lane_nveh[l] -= 1;
        c.lane_brutto[l] = c.lane_nveh[l] == 0 ? 0. : c.lane_brutto[l] - veh_length_with_gap(c, f);
        count_event(c, l, 6);
        veh_transfer_add(c, f, l);
    }
    // MSEdgeControl::executeMovements (MSEdgeControl.cpp:229-232): a lane without vehicles leaves the active list
    if (c.lane_nveh[l] == 0) {
        c.lane_actkey[l] = 0;
    }
}

// pass 2: append the arrivals at the rear
SB2_HDN void lane_handoff_integrate(const Ctx& c, int l) {
    double brutto = c.lane_brutto[l];
    // ---- MSLane::integrateNewVehicles
    int inc[16];
    int ni = 0;
    for (int v = c.lane_incoming[l]; v >= 0; v = c.veh_next_incoming[v]) {
        if (ni < 16) {
            inc[ni++] = v;
        } else {
            raise(c, ERR_LANE_ORDER);
        }
    }
    c.lane_incoming[l] = -1;
    // sort by (back position descending, numerical id descending) -- vehicle_position_sorter
    for (int i = 1; i < ni; i++) {
        const int x = inc[i];
        const double px = c.veh_pos[x] - veh_length(c, x);
        int j = i - 1;
        while (j >= 0) {
            const int y = inc[j];
            const double py = c.veh_pos[y] - veh_length(c, y);
            const bool xBeforeY = (px != py) ? (px > py) : (c.veh_numid[x] > c.veh_numid[y]);
            if (!xBeforeY) {
                break;
            }
            inc[j + 1] = y;
            j--;
        }
        inc[j + 1] = x;
    }
    for (int i = 0; i < ni; i++) {
        const int v = inc[i];
        const int oldLast = c.lane_last[l];
        c.veh_ahead[v] = oldLast;
        c.veh_behind[v] = -1;
        if (oldLast >= 0) {
            c.veh_behind[oldLast] = v;
        } else {
            c.lane_first[l] = v;
        }
        c.lane_last[l] = v;
        brutto += veh_length_with_gap(c, v);
    }
    c.lane_nveh[l] += ni;
    c.lane_brutto[l] = brutto;
    // MSEdgeControl.cpp:246-259: a lane that was empty joins the active list, in id order
    if (ni > 0 && c.lane_actkey[l] == 0) {
        c.lane_actkey[l] = lane_make_actkey(c, l, 1);
    }
    if (ni > 0) {
        c.edge_delayed[c.lane_edge[l]] = 1;   // MSEdge::markDelayed (MSLane.cpp:2608)
    }
}

// pass 4 (after the collision check, which may teleport vehicles): partial occupators were appended by executeMove; drop the
// ones that left the road, keep the rest sorted by back position (MSLane::sortPartialVehicles, MSLane.cpp:2625-2629)
SB2_HDN void lane_sort_partials(const Ctx& c, int l) {
    int np = c.lane_npartial[l];
    if (np == 0) {
        return;
    }
    if (np > MAX_PARTIAL) {
        raise(c, ERR_PARTIAL_OVERFLOW);
        np = MAX_PARTIAL;
        c.lane_npartial[l] = np;
    }
    int32_t* p = c.lane_partial + (size_t)l * MAX_PARTIAL;
    {
        // a vehicle that was teleported in this step (lane_handoff_remove) has left all lanes (MSVehicle::leaveLane resets its partial occupations)
        int keep = 0;
        for (int i = 0; i < np; i++) {
            if (c.veh_lane[p[i]] >= 0) {
                p[keep++] = p[i];
            }
        }
        if (keep != np) {
            np = keep;
            c.lane_npartial[l] = np;
        }
    }
    for (int i = 1; i < np; i++) {
        const int x = p[i];
        const double px = veh_back_pos_on_lane(c, x, l);
        int j = i - 1;
        while (j >= 0) {
            const double py = veh_back_pos_on_lane(c, p[j], l);
            // equal back positions: lower numerical id first (the reference order is then the push order)
            if (py < px || (py == px && c.veh_numid[p[j]] < c.veh_numid[x])) {
                break;
            }
            p[j + 1] = p[j];
            j--;
        }
        p[j + 1] = x;
    }
}

// MSLane::detectCollisions (non-sublane, MSLane.cpp:1826-1835, 1971-2074) with the default collision.action teleport
// (handleCollisionBetween, MSLane.cpp:2078-2211: the collider is taken off the lane and handed to MSVehicleTransfer): adjacent
// pairs front -> back, then the rear-most vehicle against the rear-most partial occupator (`*lastVeh` is myVehicles.front()).
// All pairs are tested on the queue as it is; the colliders are removed afterwards (toRemove / toTeleport sets).
SB2_HDN void lane_detect_collisions(const Ctx& c, int l) {
    const double* vt_collision_mingap_factor = c.vt_collision_mingap_factor;
    int pred = c.lane_first[l];
    if (pred < 0) {
        return;
    }
    bool any = false;
    for (int veh = c.veh_behind[pred]; veh >= 0; pred = veh, veh = c.veh_behind[veh]) {
        const double gap = (c.veh_pos[pred] - veh_length(c, pred)) - c.veh_pos[veh] - vt_collision_mingap_factor[c.veh_type[veh]] * veh_mingap(c, veh);
        if (gap < -SB2_NUMERICAL_EPS) {
            c.veh_moved[veh] = 3;   // collider: to be removed
            any = true;
        }
    }
    if (c.lane_npartial[l] > 0) {
        const int victim = c.lane_partial[(size_t)l * MAX_PARTIAL];
        const int collider = c.lane_last[l];
        const double gap = veh_back_pos_on_lane(c, victim, l) - c.veh_pos[collider] - vt_collision_mingap_factor[c.veh_type[collider]] * veh_mingap(c, collider);
        if (gap < -SB2_NUMERICAL_EPS && c.veh_moved[collider] != 3) {
            c.veh_moved[collider] = 3;
            any = true;
        }
    }
    if (!any) {
        return;
    }
    int ahead = -1;
    double brutto = c.lane_brutto[l];
    for (int x = c.lane_first[l]; x >= 0;) {
        const int next = c.veh_behind[x];
        if (c.veh_moved[x] == 3) {
            c.veh_moved[x] = 0;
            // MSLane::removeVehicle (MSLane.cpp:2870-2890)
            if (ahead >= 0) {
                c.veh_behind[ahead] = next;
            } else {
                c.lane_first[l] = next;
            }
            if (next >= 0) {
                c.veh_ahead[next] = ahead;
            } else {
                c.lane_last[l] = ahead;
            }
            c.lane_nveh[l] -= 1;
            brutto -= veh_length_with_gap(c, x);
            count_event(c, l, 4);   // MSVehicleControl::countCollision(teleport = true): a collision and a teleport
            count_event(c, l, 6);
            veh_transfer_add(c, x, l);
        } else {
            ahead = x;
        }
        x = next;
    }
    c.lane_brutto[l] = c.lane_nveh[l] == 0 ? 0. : brutto;
}

// MSTrafficLightLogic::setTrafficLightSignals(t) for phase `step` of program tl (MSTrafficLightLogic.cpp:402-414)
SB2_HD void tls_set_signals(const Ctx& c, int tl, int step, SUMOTime switchTime) {
    const uint8_t* st = c.phase_states + c.phase_state_begin[c.tl_phase_begin[tl] + step];
    for (int i = c.tl_link_begin[tl]; i < c.tl_link_begin[tl + 1]; i++) {
        const int k = c.tllink_link[i];
        const uint8_t s = st[c.tllink_index[i]];
        // MSLink::setTLState (MSLink.cpp:1325-1336)
        if (c.link_state[k] != s) {
            c.link_last_change[k] = switchTime;
        }
        c.link_state[k] = s;
        if (state_green(s)) {
            c.link_last_green[k] = s;
        }
    }
}

// static program phase clock: fires like MSTrafficLightLogic::SwitchCommand at the begin of the step
SB2_HDN void tls_advance(const Ctx& c, int tl) {
    const int p0 = c.tl_phase_begin[tl];
    const int nPhases = c.tl_phase_begin[tl + 1] - p0;
    if (nPhases == 0) {
        return;
    }
    // a phase of duration 0 would re-fire within the same step in the reference's event queue
    int guard = 0;
    while (c.tl_next_switch[tl] <= c.t && guard++ < 16) {
        // MSSimpleTrafficLightLogic::trySwitch: step = (step + 1) % phases ; return duration
        int step = c.tl_cur_phase[tl] + 1;
        if (step >= nPhases) {
            step = 0;
        }
        c.tl_cur_phase[tl] = step;
        const SUMOTime switchTime = c.tl_next_switch[tl];
        c.tl_next_switch[tl] = switchTime + c.phase_dur[p0 + step];
        tls_set_signals(c, tl, step, switchTime);
    }
}

// libsumo::TrafficLight::setPhase (src/libsumo/TrafficLight.cpp:823-832) -> MSSimpleTrafficLightLogic::changeStepAndDuration
// (MSSimpleTrafficLightLogic.cpp:294-309) between two steps: the pending switch is rescheduled to now + the phase's duration;
// if the phase changes, the signals are set at once with the current time as their change time
SB2_HD void tls_set_phase(const Ctx& c, int tl, int step) {
    const SUMOTime now = c.t;
    c.tl_next_switch[tl] = now + c.phase_dur[c.tl_phase_begin[tl] + step];
    if (step != c.tl_cur_phase[tl]) {
        c.tl_cur_phase[tl] = step;
        tls_set_signals(c, tl, step, now);
    }
}

}  // namespace sb2
