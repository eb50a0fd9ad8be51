/* sb2_e1.h -- E1 induction loops as move reminders of the vehicles.
 *
 * Restates src/microsim/output/MSInductLoop.cpp:119-146 (notifyEnter), 149-209 (notifyMove), 212-230 (notifyLeave), 561-565
 * (VehicleData), src/microsim/cfmodels/MSCFModel.cpp:691-724 (passingTime, Euler update) and the reminder bookkeeping of the
 * vehicle: src/microsim/MSVehicle.cpp:1157-1183 (workOnMoveReminders), 1205-1223 (adaptLaneEntering2MoveReminder), 5565-5591
 * (enterLaneAtMove), 5813-5822 (enterLaneAtInsertion), 5862-5880 (leaveLane), src/microsim/MSBaseVehicle.cpp:1069-1098
 * (activateReminders).
 *
 * A vehicle carries its active reminders (detector, offset = length of the lanes it left since the detector's lane, entry time if
 * it is on the detector) in MAX_E1_REMINDERS slots; every passage becomes one E1Event record in a device list (atomic append) that
 * sumob200_e1_interval reduces per detector on the host in a fixed order (detector, leave time, vehicle) -- the reference adds
 * them in notification order; the sums are printed with two decimals.
 * Not modelled: vTypes / nextEdges filters, persons, the override of TraCI, mesoscopic mode. */
#pragma once
#include "sb2_types.h"

namespace sb2 {

// reminder slot i of vehicle v (slot-major: re-slotting permutes MAX_E1_REMINDERS rows of nV)
#define E1_AT(i) ((size_t)(i) * c.nV + v)

// MSCFModel::passingTime, gSemiImplicitEulerUpdate (MSCFModel.cpp:691-724)
SB2_HD double cf_passingTime(const Ctx& c, double lastPos, double passedPos, double currentPos, double currentSpeed) {
    if (passedPos > currentPos || passedPos < lastPos) {
        const double lastCoveredDist = currentPos - lastPos;
        return passedPos > currentPos ? c.TS * (passedPos - lastPos) / lastCoveredDist : c.TS * (currentPos - passedPos) / lastCoveredDist;
    }
    const double distanceOldToPassed = passedPos - lastPos;
    if (currentSpeed == 0) {
        return c.TS;
    }
    const double t = distanceOldToPassed / currentSpeed;
    return MIN2(c.TS, MAX2(0., t));
}

SB2_HD double e1_simtime(const Ctx& c) { return (double)c.t / 1000.; }   // SIMTIME = STEPS2TIME(getCurrentTimeStep())

// myVehicleDataCont.push_back(VehicleData(veh, entryTime, leaveTime, leftEarly, ...))
// (edge partition: a detector is reported by the rank that owns its lane; the vehicles of the ghost column are executed redundantly)
SB2_HD bool e1_owned(const Ctx& c, int det) { return lane_auth(c, c.e1_lane[det]); }

SB2_HD void e1_push_event(const Ctx& c, int det, int v, double entry, double leave, bool leftEarly) {
    if (!e1_owned(c, det)) {
        return;
    }
#if defined(__CUDA_ARCH__)
    const int at = atomicAdd(&c.counters[12], 1);
#else
    const int at = c.counters[12]++;
#endif
    if (at >= c.e1_event_cap) {
        raise(c, ERR_DETECTOR_OVERFLOW);
        return;
    }
    E1Event ev;
    ev.det = det; ev.veh = c.veh_ext[v]; ev.entry = entry; ev.leave = leave;
    ev.length = c.vt_length[c.veh_type[v]];
    ev.leftEarly = leftEarly ? 1 : 0; ev.pad = 0;
    c.e1_events[at] = ev;
}

SB2_HD void e1_count_entered(const Ctx& c, int det) {
    if (!e1_owned(c, det)) {
        return;
    }
#if defined(__CUDA_ARCH__)
    atomicAdd(&c.e1_entered[det], 1);
#else
    c.e1_entered[det]++;
#endif
}

// edge partition: a vehicle record received for a lane this rank does not simulate -- whatever reminders the slot still holds are
// from an earlier visit of the vehicle to this rank's strip
SB2_HD void veh_e1_forget(const Ctx& c, int v) {
    if (!SB2_OBSERVERS || c.nE1 == 0) {
        return;
    }
    for (int i = 0; i < MAX_E1_REMINDERS; i++) {
        c.veh_e1_det[E1_AT(i)] = -1;
        c.veh_e1_entry[E1_AT(i)] = E1_NOT_ON;
    }
}

// MSVehicle::leaveLane(reason != NOTIFICATION_JUNCTION): notifyLeave of every reminder -> a vehicle on a detector leaves it early
// (arrival, teleport, collision); no reminder survives
SB2_HD void veh_e1_leave_net(const Ctx& c, int v) {
    if (!SB2_OBSERVERS || c.nE1 == 0) {
        return;
    }
    for (int i = 0; i < MAX_E1_REMINDERS; i++) {
        const int det = c.veh_e1_det[E1_AT(i)];
        if (det >= 0) {
            const double entry = c.veh_e1_entry[E1_AT(i)];
            if (entry != E1_NOT_ON) {
                e1_push_event(c, det, v, entry, e1_simtime(c) + c.TS, true);
            }
            c.veh_e1_det[E1_AT(i)] = -1;
            c.veh_e1_entry[E1_AT(i)] = E1_NOT_ON;
        }
    }
}

// the detectors of `lane` become reminders with offset 0 (addReminder); `enter`: notifyEnter with a reason other than
// NOTIFICATION_JUNCTION (departure) -- dropped if the vehicle's back is beyond the detector, on the detector if its front is
SB2_HD void veh_e1_add_lane(const Ctx& c, int v, int lane, bool enter, double pos, double backPos) {
    if (!SB2_OBSERVERS || c.nE1 == 0) {
        return;
    }
    for (int k = c.lane_e1_begin[lane]; k < c.lane_e1_begin[lane + 1]; k++) {
        const int det = c.e1_of_lane[k];
        double entry = E1_NOT_ON;
        if (enter) {
            if (backPos >= c.e1_pos[det]) {
                continue;
            }
            if (pos >= c.e1_pos[det]) {
                entry = e1_simtime(c);
                e1_count_entered(c, det);
            }
        }
        int slot = -1;
        for (int i = 0; i < MAX_E1_REMINDERS; i++) {
            if (c.veh_e1_det[E1_AT(i)] < 0) {
                slot = i;
                break;
            }
        }
        if (slot < 0) {
            raise(c, ERR_DETECTOR_OVERFLOW);
            continue;
        }
        c.veh_e1_det[E1_AT(slot)] = det;
        c.veh_e1_off[E1_AT(slot)] = 0.;
        c.veh_e1_entry[E1_AT(slot)] = entry;
    }
}

// adaptLaneEntering2MoveReminder: rem.second += oldLaneLength, then the entered lane's reminders (notifyEnter(JUNCTION) keeps them)
SB2_HD void veh_e1_enter_at_move(const Ctx& c, int v, int oldLane, int newLane) {
    if (!SB2_OBSERVERS || c.nE1 == 0) {
        return;
    }
    const double oldLaneLength = c.lane_length[oldLane];
    for (int i = 0; i < MAX_E1_REMINDERS; i++) {
        if (c.veh_e1_det[E1_AT(i)] >= 0) {
            c.veh_e1_off[E1_AT(i)] += oldLaneLength;
        }
    }
    veh_e1_add_lane(c, v, newLane, false, 0., 0.);
}

// workOnMoveReminders(oldPos, newPos, newSpeed) -> MSInductLoop::notifyMove for every reminder
SB2_HD void veh_e1_move(const Ctx& c, int v, double oldPos0, double newPos0, double newSpeed0) {
    if (!SB2_OBSERVERS || c.nE1 == 0) {
        return;
    }
    const double newSpeed = MAX2(0., newSpeed0);
    const double length = c.vt_length[c.veh_type[v]];
    for (int i = 0; i < MAX_E1_REMINDERS; i++) {
        const int det = c.veh_e1_det[E1_AT(i)];
        if (det < 0) {
            continue;
        }
        const double off = c.veh_e1_off[E1_AT(i)];
        const double oldPos = oldPos0 + off, newPos = newPos0 + off;
        const double myPosition = c.e1_pos[det], myEndPosition = c.e1_end[det];
        if (newPos < myPosition) {
            continue;   // detector not reached yet
        }
        double entry = c.veh_e1_entry[E1_AT(i)];
        if (newPos >= myPosition && oldPos < myPosition) {
            // entered the detector by move
            const double timeBeforeEnter = cf_passingTime(c, oldPos, myPosition, newPos, newSpeed);
            entry = e1_simtime(c) + timeBeforeEnter;
            c.veh_e1_entry[E1_AT(i)] = entry;
            e1_count_entered(c, det);
        }
        const double oldBackPos = oldPos - length;
        const double newBackPos = newPos - length;
        if (newBackPos > myEndPosition) {
            // vehicle passed the detector
            if (oldBackPos <= myEndPosition && entry != E1_NOT_ON) {
                const double leaveTime = e1_simtime(c) + cf_passingTime(c, oldBackPos, myEndPosition, newBackPos, newSpeed);
                e1_push_event(c, det, v, entry, leaveTime, false);
            }
            c.veh_e1_det[E1_AT(i)] = -1;
            c.veh_e1_entry[E1_AT(i)] = E1_NOT_ON;
        }
    }
}

// the vehicles that are on a detector right now (myVehiclesOnDet), as pseudo events with leave == E1_STILL_ON: they add to the
// occupancy of the interval that is being written (MSInductLoop.cpp:479-481)
SB2_HD void veh_e1_collect_on_detector(const Ctx& c, int v) {
    for (int i = 0; i < MAX_E1_REMINDERS; i++) {
        const int det = c.veh_e1_det[E1_AT(i)];
        if (det >= 0 && c.veh_e1_entry[E1_AT(i)] != E1_NOT_ON) {
            e1_push_event(c, det, v, c.veh_e1_entry[E1_AT(i)], E1_STILL_ON, false);
        }
    }
}

#undef E1_AT

}  // namespace sb2
