/* sb2_e2.h -- E2 lane area detectors on one lane as move reminders of the vehicles.
 *
 * Restates src/microsim/output/MSE2Collector.cpp:590-680 (notifyMove), 683-737 (notifyLeave), 740-831 (notifyEnter), 835-848
 * (makeVehicleInfo), 1056-1100 (makeMoveNotification), 1268-1325 (calculateTimeLossAndTimeOnDetector) on the vehicle's side and
 * 874-992 (detectorUpdate), 995-1016 (aggregateOutputValues), 1019-1051 (integrateMoveNotification), 1103-1155 (buildJam),
 * 1158-1211 (checkJam), 1214-1265 (processJams), 1333-1422 (writeXMLOutput), 1425-1451 (reset) on the detector's side.
 *
 * A vehicle carries its subscriptions (detector, rem->second, hasEntered) in MAX_E2_REMINDERS slots and pushes one E2Note per
 * detector and step while it is on / before the detector; after the step one thread per detector sorts its notes by the distance to
 * the detector's end (the reference's std::sort key), walks them in that order -- the fp64 sums are added in the reference's order,
 * so the accumulators are the reference's bit for bit -- keeps the halting durations (the reference's two std::maps by vehicle id
 * are the notes of the previous step here) and builds the jams.  e2_detector_readout = writeXMLOutput + reset.
 * Not modelled: detectors over several lanes, vTypes / nextEdges filters, persons, detectors coupled to a traffic light. */
#pragma once
#include "sb2_e1.h"

namespace sb2 {

#define E2_AT(i) ((size_t)(i) * c.nV + v)

// (edge partition: a detector is reported by the rank that owns its lane)
SB2_HD bool e2_owned(const Ctx& c, int det) { return lane_auth(c, c.e2_lane[det]); }

SB2_HD void e2_count(const Ctx& c, int det, int k) {
    if (e2_owned(c, det)) {
        atomic_inc(&c.e2_cnt[(size_t)k * c.nE2 + det]);
    }
}

// myMoveNotifications.push_back(makeMoveNotification(...))
SB2_HD void e2_push_note(const Ctx& c, int det, int v, double dist, double timeOn, double lenOn, double timeLoss, double speed) {
    if (!e2_owned(c, det)) {
        return;
    }
    const int buf = c.cur;
#if defined(__CUDA_ARCH__)
    const int at = atomicAdd(&c.counters[13 + buf], 1);
#else
    const int at = c.counters[13 + buf]++;
#endif
    if (at >= c.e2_note_cap) {
        raise(c, ERR_DETECTOR_OVERFLOW);
        return;
    }
    E2Note n;
    n.veh = c.veh_ext[v]; n.dist = dist; n.timeOn = timeOn; n.lenOn = lenOn; n.timeLoss = timeLoss; n.speed = speed; n.halt = 0; n.ihalt = 0;
#if defined(__CUDA_ARCH__)
    n.next = atomicExch(&c.e2_head[buf][det], at);
#else
    n.next = c.e2_head[buf][det];
    c.e2_head[buf][det] = at;
#endif
    c.e2_notes[buf][at] = n;
}

// edge partition: see veh_e1_forget
SB2_HD void veh_e2_forget(const Ctx& c, int v) {
    if (!SB2_OBSERVERS || c.nE2 == 0) {
        return;
    }
    for (int i = 0; i < MAX_E2_REMINDERS; i++) {
        c.veh_e2_det[E2_AT(i)] = -1;
    }
}

// notifyLeave with a reason other than NOTIFICATION_JUNCTION (arrival, teleport): the VehicleInfo is erased at once
SB2_HD void veh_e2_leave_net(const Ctx& c, int v) {
    if (!SB2_OBSERVERS || c.nE2 == 0) {
        return;
    }
    for (int i = 0; i < MAX_E2_REMINDERS; i++) {
        const int det = c.veh_e2_det[E2_AT(i)];
        if (det >= 0) {
            if (c.veh_e2_entered[E2_AT(i)]) {
                e2_count(c, det, E2C_LEFT);
            }
            c.veh_e2_det[E2_AT(i)] = -1;
        }
    }
}

// notifyEnter for the detectors of `lane`; `enter`: reason other than NOTIFICATION_JUNCTION (departure, end of a teleport) -- no
// subscription if the vehicle's back is already behind the detector's end
SB2_HD void veh_e2_add_lane(const Ctx& c, int v, int lane, bool enter, double backPos) {
    if (!SB2_OBSERVERS || c.nE2 == 0) {
        return;
    }
    for (int k = c.lane_e2_begin[lane]; k < c.lane_e2_begin[lane + 1]; k++) {
        const int det = c.e2_of_lane[k];
        if (enter && c.e2_start[det] + c.e2_length[det] <= backPos) {
            continue;
        }
        int slot = -1;
        bool known = false;
        for (int i = 0; i < MAX_E2_REMINDERS; i++) {
            const int d = c.veh_e2_det[E2_AT(i)];
            if (d == det) {
                known = true;   // "already known. No new VehicleInfo is created"
            } else if (d < 0 && slot < 0) {
                slot = i;
            }
        }
        if (known) {
            continue;
        }
        if (slot < 0) {
            raise(c, ERR_DETECTOR_OVERFLOW);
            continue;
        }
        c.veh_e2_det[E2_AT(slot)] = det;
        c.veh_e2_off[E2_AT(slot)] = 0.;
        c.veh_e2_entered[E2_AT(slot)] = 0;
    }
}

// adaptLaneEntering2MoveReminder + the entered lane's reminders
SB2_HD void veh_e2_enter_at_move(const Ctx& c, int v, int oldLane, int newLane) {
    if (!SB2_OBSERVERS || c.nE2 == 0) {
        return;
    }
    const double oldLaneLength = c.lane_length[oldLane];
    for (int i = 0; i < MAX_E2_REMINDERS; i++) {
        if (c.veh_e2_det[E2_AT(i)] >= 0) {
            c.veh_e2_off[E2_AT(i)] += oldLaneLength;
        }
    }
    veh_e2_add_lane(c, v, newLane, false, 0.);
}

// workOnMoveReminders(oldPos, newPos, newSpeed) -> MSE2Collector::notifyMove of every subscription; vmaxLane =
// veh.getLane()->getVehicleMaxSpeed(&veh) of the lane the vehicle is on after the move
SB2_HD void veh_e2_move(const Ctx& c, int v, double oldPos0, double newPos0, double speed, double vmaxLane) {
    if (!SB2_OBSERVERS || c.nE2 == 0) {
        return;
    }
    const double TS = c.TS;
    const double length = c.vt_length[c.veh_type[v]];
    for (int i = 0; i < MAX_E2_REMINDERS; i++) {
        const int det = c.veh_e2_det[E2_AT(i)];
        if (det < 0) {
            continue;
        }
        const double off = c.veh_e2_off[E2_AT(i)];
        const double oldPos = oldPos0 + off, newPos = newPos0 + off;
        const double detLength = c.e2_length[det];
        const double entryOffset = -c.e2_start[det];                 // myOffsets[0]
        double exitOffset = entryOffset - detLength;                 // makeVehicleInfo
        if (off > 0.) {
            // notifyLeave(NOTIFICATION_JUNCTION) into a lane that is not the detector's
            exitOffset = MAX2(exitOffset, entryOffset - entryOffset - c.lane_length[c.e2_lane[det]]);
        }
        const double relPos = entryOffset + newPos;
        if (relPos <= 0) {
            continue;   // detector not yet reached
        }
        if (!c.veh_e2_entered[E2_AT(i)]) {
            c.veh_e2_entered[E2_AT(i)] = 1;
            e2_count(c, det, E2C_ENTERED);
            e2_count(c, det, E2C_SEEN);
        }
        const bool vehPassedDetectorEnd = -exitOffset <= newPos - length;
        const bool vehicleEnteredLaneAfterDetector = vehPassedDetectorEnd && (-exitOffset <= oldPos - length);
        if (!vehicleEnteredLaneAfterDetector) {
            // ---- calculateTimeLossAndTimeOnDetector
            double timeOnDetector, timeLoss;
            if (oldPos == newPos) {
                timeLoss = TS;
                timeOnDetector = TS;
            } else {
                const double entryPos = MAX2(-entryOffset, 0.);
                double entryTime = 0;
                if (oldPos < entryPos) {
                    entryTime = cf_passingTime(c, oldPos, entryPos, newPos, speed);
                }
                const double entrySpeed = (newPos - oldPos) / TS;   // speedAfterTime, Euler: DIST2SPEED(dist)
                const double exitPos = MIN2(newPos, -exitOffset + length);
                double exitTime;
                if (exitPos == newPos) {
                    exitTime = TS;
                } else {
                    exitTime = cf_passingTime(c, oldPos, exitPos, newPos, speed);
                }
                const double exitSpeed = (newPos - oldPos) / TS;
                const double vmax = MAX2(vmaxLane, SB2_NUMERICAL_EPS);
                timeOnDetector = exitTime - entryTime;
                timeLoss = MAX2(0., timeOnDetector * (vmax - (entrySpeed + exitSpeed) / 2) / vmax);
            }
            // ---- makeMoveNotification
            double lengthOnDetector = MAX2(MIN2(length, newPos + entryOffset), 0.);
            const double distToExit = -exitOffset - newPos;
            lengthOnDetector = MAX2(0., lengthOnDetector + MIN2(0., distToExit));
            e2_push_note(c, det, v, detLength - (entryOffset + newPos), timeOnDetector, lengthOnDetector, timeLoss, MAX2(0., speed));
        }
        if (vehPassedDetectorEnd) {
            e2_count(c, det, E2C_LEFT);   // myLeftVehicles -> erased in detectorUpdate of this step
            c.veh_e2_det[E2_AT(i)] = -1;
        }
    }
}

#undef E2_AT

// ---------------------------------------------------------------- the detector's side: one thread per detector, after the step

#define E2D(k) c.e2_acc_d[(size_t)(k) * c.nE2 + d]
#define E2I(k) c.e2_acc_i[(size_t)(k) * c.nE2 + d]

// MSE2Collector::detectorUpdate
SB2_HDN void e2_detector_update(const Ctx& c, int d) {
    const int buf = c.cur, prev = c.cur ^ 1;
    E2Note* notes = c.e2_notes[buf];
    const E2Note* old = c.e2_notes[prev];
    // std::sort(myMoveNotifications, compareMoveNotification): ascending distance to the detector's end (ties: vehicle index)
    int sorted = -1;
    int numVehicles = 0;
    for (int n = c.e2_head[buf][d]; n >= 0;) {
        const int next = notes[n].next;
        int before = -1;
        for (int m = sorted; m >= 0; m = notes[m].next) {
            if (notes[m].dist > notes[n].dist || (notes[m].dist == notes[n].dist && notes[m].veh > notes[n].veh)) {
                break;
            }
            before = m;
        }
        if (before < 0) {
            notes[n].next = sorted;
            sorted = n;
        } else {
            notes[n].next = notes[before].next;
            notes[before].next = n;
        }
        numVehicles++;
        n = next;
    }
    c.e2_head[buf][d] = sorted;
    double currentMeanLength = 0;
    double currentMaxJamLengthInMeters = 0;
    int64_t currentMaxJamLengthInVehicles = 0;
    const double jamSpeed = c.e2_jam_speed[d], jamDist = c.e2_jam_dist[d];
    const int64_t jamTime = c.e2_jam_time[d];
    int jamFirst = -1, jamLast = -1, jamFirstIdx = 0, jamLastIdx = 0;   // currentJam
    int idx = 0;
    for (int n = sorted; n >= 0; n = notes[n].next, idx++) {
        E2Note& mni = notes[n];
        // integrateMoveNotification
        E2D(E2D_SAMPLES) += mni.timeOn;
        E2D(E2D_TIMELOSS) += mni.timeLoss;
        E2D(E2D_SPEEDSUM) += mni.speed * mni.timeOn;
        currentMeanLength += mni.lenOn;
        // checkJam
        int64_t wasHalt = 0, wasIHalt = 0;
        for (int m = c.e2_head[prev][d]; m >= 0; m = old[m].next) {
            if (old[m].veh == mni.veh) {
                wasHalt = old[m].halt;
                wasIHalt = old[m].ihalt;
                break;
            }
        }
        bool isInJam = false;
        if (mni.speed < jamSpeed) {
            if (wasHalt > 0) {
                mni.halt = wasHalt + c.DELTA_T;
                mni.ihalt = wasIHalt + c.DELTA_T;
            } else {
                mni.halt = c.DELTA_T;
                mni.ihalt = c.DELTA_T;
                E2I(E2I_STARTEDHALTS) += 1;
            }
            if (mni.halt > jamTime) {
                isInJam = true;
            }
        } else if (wasHalt > 0) {
            // not standing anymore: myPastStandingDurations / myPastIntervalStandingDurations
            E2I(E2I_PAST_SUM) += wasHalt;
            E2I(E2I_PAST_MAX) = wasHalt > E2I(E2I_PAST_MAX) ? wasHalt : E2I(E2I_PAST_MAX);
            E2I(E2I_PAST_N) += 1;
            E2I(E2I_IPAST_SUM) += wasIHalt;
            E2I(E2I_IPAST_MAX) = wasIHalt > E2I(E2I_IPAST_MAX) ? wasIHalt : E2I(E2I_IPAST_MAX);
            E2I(E2I_IPAST_N) += 1;
        }
        // buildJam (+ processJams for a jam that is closed)
        bool close = false;
        if (isInJam) {
            if (jamFirst >= 0 && notes[jamLast].dist - mni.dist > jamDist) {
                close = true;
            }
        } else if (jamFirst >= 0) {
            close = true;
        }
        if (close) {
            const double jamLengthInMeters = MAX2(notes[jamLast].dist, 0.) - MAX2(notes[jamFirst].dist, 0.) + notes[jamLast].lenOn;
            const int64_t jamLengthInVehicles = jamLastIdx - jamFirstIdx + 1;
            currentMaxJamLengthInMeters = MAX2(currentMaxJamLengthInMeters, jamLengthInMeters);
            currentMaxJamLengthInVehicles = jamLengthInVehicles > currentMaxJamLengthInVehicles ? jamLengthInVehicles : currentMaxJamLengthInVehicles;
            E2D(E2D_JAMSUM_M) += jamLengthInMeters;
            E2I(E2I_JAMSUM_V) += jamLengthInVehicles;
            jamFirst = -1;
        }
        if (isInJam) {
            if (jamFirst < 0) {
                jamFirst = n;
                jamFirstIdx = idx;
            }
            jamLast = n;
            jamLastIdx = idx;
        }
    }
    if (jamFirst >= 0) {
        const double jamLengthInMeters = MAX2(notes[jamLast].dist, 0.) - MAX2(notes[jamFirst].dist, 0.) + notes[jamLast].lenOn;
        const int64_t jamLengthInVehicles = jamLastIdx - jamFirstIdx + 1;
        currentMaxJamLengthInMeters = MAX2(currentMaxJamLengthInMeters, jamLengthInMeters);
        currentMaxJamLengthInVehicles = jamLengthInVehicles > currentMaxJamLengthInVehicles ? jamLengthInVehicles : currentMaxJamLengthInVehicles;
        E2D(E2D_JAMSUM_M) += jamLengthInMeters;
        E2I(E2I_JAMSUM_V) += jamLengthInVehicles;
    }
    // aggregateOutputValues
    E2I(E2I_TIMESAMPLES) += 1;
    const double currentOccupancy = currentMeanLength / c.e2_length[d] * (double)100.;
    E2D(E2D_OCCSUM) += currentOccupancy;
    E2D(E2D_MAXOCC) = MAX2(E2D(E2D_MAXOCC), currentOccupancy);
    E2I(E2I_MEANMAXJAM_V) += currentMaxJamLengthInVehicles;
    E2D(E2D_MEANMAXJAM_M) += currentMaxJamLengthInMeters;
    E2I(E2I_MAXJAM_V) = currentMaxJamLengthInVehicles > E2I(E2I_MAXJAM_V) ? currentMaxJamLengthInVehicles : E2I(E2I_MAXJAM_V);
    E2D(E2D_MAXJAM_M) = MAX2(E2D(E2D_MAXJAM_M), currentMaxJamLengthInMeters);
    E2I(E2I_MEANVEHNUM) += numVehicles;
    E2I(E2I_MAXVEHNUM) = numVehicles > E2I(E2I_MAXVEHNUM) ? numVehicles : E2I(E2I_MAXVEHNUM);
    // the other buffer takes the notes of the next step
    c.e2_head[prev][d] = -1;
    if (d == 0) {
        c.counters[13 + prev] = 0;
    }
}

// MSE2Collector::writeXMLOutput + reset: the attributes of one <interval> (times in seconds), then the interval's accumulators start
// over.  `buf` = the buffer of the last step that was processed (its notes carry the halting durations of the vehicles on the detector).
SB2_HDN void e2_detector_readout(const Ctx& c, int d, int buf, E2Result* out) {
    E2Note* notes = c.e2_notes[buf];
    int32_t* cnt = c.e2_cnt;
    const int nEntered = cnt[(size_t)E2C_ENTERED * c.nE2 + d], nLeft = cnt[(size_t)E2C_LEFT * c.nE2 + d], nSeen = cnt[(size_t)E2C_SEEN * c.nE2 + d];
    const int64_t timeSamples = E2I(E2I_TIMESAMPLES);
    E2Result r;
    r.sampledSeconds = E2D(E2D_SAMPLES);
    r.nVehEntered = nEntered;
    r.nVehLeft = nLeft;
    r.nVehSeen = nSeen;
    r.meanSpeed = E2D(E2D_SAMPLES) != 0 ? E2D(E2D_SPEEDSUM) / E2D(E2D_SAMPLES) : -1;
    r.meanTimeLoss = nSeen != 0 ? E2D(E2D_TIMELOSS) / nSeen : -1;
    r.meanOccupancy = timeSamples != 0 ? E2D(E2D_OCCSUM) / (double)timeSamples : 0;
    r.maxOccupancy = E2D(E2D_MAXOCC);
    r.meanMaxJamLengthInVehicles = timeSamples != 0 ? (double)E2I(E2I_MEANMAXJAM_V) / (double)timeSamples : 0;
    r.meanMaxJamLengthInMeters = timeSamples != 0 ? E2D(E2D_MEANMAXJAM_M) / (double)timeSamples : 0;
    r.maxJamLengthInVehicles = (int32_t)E2I(E2I_MAXJAM_V);
    r.maxJamLengthInMeters = E2D(E2D_MAXJAM_M);
    r.jamLengthInVehiclesSum = (int32_t)E2I(E2I_JAMSUM_V);
    r.jamLengthInMetersSum = E2D(E2D_JAMSUM_M);
    int64_t haltingDurationSum = E2I(E2I_PAST_SUM), maxHaltingDuration = E2I(E2I_PAST_MAX), haltingNo = E2I(E2I_PAST_N);
    int64_t intervalHaltingDurationSum = E2I(E2I_IPAST_SUM), intervalMaxHaltingDuration = E2I(E2I_IPAST_MAX), intervalHaltingNo = E2I(E2I_IPAST_N);
    for (int n = c.e2_head[buf][d]; n >= 0; n = notes[n].next) {
        if (notes[n].halt > 0) {   // myHaltingVehicleDurations / myIntervalHaltingVehicleDurations
            haltingDurationSum += notes[n].halt;
            maxHaltingDuration = notes[n].halt > maxHaltingDuration ? notes[n].halt : maxHaltingDuration;
            haltingNo++;
            intervalHaltingDurationSum += notes[n].ihalt;
            intervalMaxHaltingDuration = notes[n].ihalt > intervalMaxHaltingDuration ? notes[n].ihalt : intervalMaxHaltingDuration;
            intervalHaltingNo++;
            notes[n].ihalt = 0;    // reset()
        }
    }
    const int64_t meanHaltingDuration = haltingNo != 0 ? haltingDurationSum / haltingNo : 0;
    const int64_t intervalMeanHaltingDuration = intervalHaltingNo != 0 ? intervalHaltingDurationSum / intervalHaltingNo : 0;
    r.meanHaltingDuration = STEPS2TIME(meanHaltingDuration);
    r.maxHaltingDuration = STEPS2TIME(maxHaltingDuration);
    r.haltingDurationSum = STEPS2TIME(haltingDurationSum);
    r.meanIntervalHaltingDuration = STEPS2TIME(intervalMeanHaltingDuration);
    r.maxIntervalHaltingDuration = STEPS2TIME(intervalMaxHaltingDuration);
    r.intervalHaltingDurationSum = STEPS2TIME(intervalHaltingDurationSum);
    r.startedHalts = (int32_t)E2I(E2I_STARTEDHALTS);
    r.meanVehicleNumber = timeSamples != 0 ? (double)E2I(E2I_MEANVEHNUM) / (double)timeSamples : 0;
    r.maxVehicleNumber = (int32_t)E2I(E2I_MAXVEHNUM);
    r.pad = 0;
    out[d] = r;
    // reset()
    for (int k = 0; k < E2D_N; k++) {
        E2D(k) = 0;
    }
    for (int k = 0; k < E2I_N; k++) {
        E2I(k) = 0;
    }
    cnt[(size_t)E2C_ENTERED * c.nE2 + d] = 0;
    cnt[(size_t)E2C_SEEN * c.nE2 + d] = nSeen - nLeft;
    cnt[(size_t)E2C_LEFT * c.nE2 + d] = 0;
}

#undef E2D
#undef E2I

}  // namespace sb2
