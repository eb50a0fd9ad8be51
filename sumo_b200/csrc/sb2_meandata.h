/* sb2_meandata.h -- edgeData / laneData (MSMeanData_Net) as move reminders of the vehicles.
 *
 * Restates src/microsim/output/MSMeanData.cpp:87-253 (MeanDataValues::notifyMove: the part of the step a vehicle spends on a lane,
 * with the entry / exit instants interpolated by MSCFModel::passingTime and the occupied length integrated), and
 * src/microsim/output/MSMeanData_Net.cpp:131-189 (notifyMoveInternal), 192-221 (notifyLeave), 224-246 (notifyEnter).
 *
 * The reference hangs one MSLaneMeanDataValues on every (non-internal) lane; a vehicle keeps it as a move reminder from entering
 * the lane until its back has left it.  Here the lanes a vehicle touches in a step are known in executeMove (the lanes its body
 * was on + the lanes it entered), so no per-vehicle reminder list is stored: md_notifyMove is called for each of them with the
 * positions shifted by the reference's rem->second (the lengths of the lanes left since, summed in the same order).
 * Per lane the engine accumulates the raw sums of MSLaneMeanDataValues (atomics: the order of additions inside a step is not
 * the reference's; the values are printed with two decimals); the derived attributes (density, traveltime, ...) stay with the
 * reference's MSLaneMeanDataValues::write.  Defaults of <edgeData>: no internal lanes, all vehicle types, haltingSpeedThreshold
 * given by the scenario (md_halt_speed), no relative threshold, trackVehicles off. */
#pragma once
#include "sb2_e1.h"
#include "sb2_veh.h"

namespace sb2 {


// (edge partition: a rank accumulates its OWN lanes only -- it executes every vehicle that touches them, own or ghost; the lanes of
// its ghost column are the neighbour's to report)
SB2_HD bool md_lane_measured(const Ctx& c, int lane) { return SB2_OBSERVERS && c.md_on != 0 && !lane_internal(c, lane) && lane_auth(c, lane); }

SB2_HD void md_add(const Ctx& c, int lane, int k, double x) {
    double* p = &c.md_sums[(size_t)k * c.nL + lane];
#if defined(__CUDA_ARCH__)
    atomicAdd(p, x);
#else
    *p += x;
#endif
}

// notifyEnter / notifyLeave counters of the lane the vehicle is on (getLane() == veh.getLane())
SB2_HD void md_count(const Ctx& c, int lane, int k) {
    if (!md_lane_measured(c, lane)) {
        return;
    }
    atomic_inc(&c.md_counts[(size_t)k * c.nL + lane]);
}

// MeanDataValues::notifyMove + MSLaneMeanDataValues::notifyMoveInternal for the reminder of `lane`; oldPos / newPos relative to the
// lane's start; vmax = the speed limit of the lane the vehicle is on NOW for this vehicle (veh.getLane()->getVehicleMaxSpeed(&veh))
SB2_HD void md_notifyMove(const Ctx& c, int v, int lane, double oldPos, double newPos, double newSpeed, double vmax) {
    if (!md_lane_measured(c, lane)) {
        return;
    }
    const double TS = c.TS;
    const double myLaneLength = c.lane_length[lane];
    const double vehLength = veh_length(c, v);
    double enterSpeed = newSpeed;   // gSemiImplicitEulerUpdate
    double leaveSpeed = newSpeed, leaveSpeedFront = newSpeed;
    double timeOnLane = TS;
    double frontOnLane = oldPos > myLaneLength ? 0. : TS;
    double timeBeforeEnter = 0.;
    double timeBeforeEnterBack = 0.;
    double timeBeforeLeaveFront = newPos <= myLaneLength ? TS : 0.;
    double timeBeforeLeave = TS;
    if (oldPos < 0 && newPos >= 0) {
        timeBeforeEnter = cf_passingTime(c, oldPos, 0, newPos, newSpeed);
        timeOnLane = TS - timeBeforeEnter;
        frontOnLane = timeOnLane;
        enterSpeed = (newPos - oldPos) / TS;   // speedAfterTime, Euler: DIST2SPEED(dist)
    }
    const double oldBackPos = oldPos - vehLength;
    const double newBackPos = newPos - vehLength;
    if (oldBackPos < 0. && newBackPos > 0.) {
        timeBeforeEnterBack = cf_passingTime(c, oldBackPos, 0., newBackPos, newSpeed);
    } else if (newBackPos <= 0) {
        timeBeforeEnterBack = TS;
    } else {
        timeBeforeEnterBack = 0.;
    }
    if (newBackPos > myLaneLength && oldBackPos <= myLaneLength) {
        timeBeforeLeave = cf_passingTime(c, oldBackPos, myLaneLength, newBackPos, newSpeed);
        const double timeAfterLeave = TS - timeBeforeLeave;
        timeOnLane -= timeAfterLeave;
        leaveSpeed = (newPos - oldPos) / TS;
        if (fabs(timeOnLane) < SB2_NUMERICAL_EPS) {
            timeOnLane = 0.;
        }
    }
    if (newPos > myLaneLength && oldPos <= myLaneLength) {
        timeBeforeLeaveFront = cf_passingTime(c, oldPos, myLaneLength, newPos, newSpeed);
        const double timeAfterLeave = TS - timeBeforeLeaveFront;
        frontOnLane -= timeAfterLeave;
        if (fabs(frontOnLane) < SB2_NUMERICAL_EPS) {
            frontOnLane = 0.;
        }
        leaveSpeedFront = (newPos - oldPos) / TS;
    }
    (void)leaveSpeedFront;
    if (timeOnLane <= 0) {
        return;
    }
    // average vehicle length on the lane in the step
    const double lengthOnLaneAtStepStart = MAX2(0., MIN2(MIN2(myLaneLength, vehLength), MIN2(vehLength - (oldPos - myLaneLength), oldPos)));
    const double lengthOnLaneAtStepEnd = MAX2(0., MIN2(MIN2(myLaneLength, vehLength), MIN2(vehLength - (newPos - myLaneLength), newPos)));
    double integratedLengthOnLane = 0.;
    if (timeBeforeEnterBack < timeBeforeLeaveFront) {
        const double lengthOnLaneAtBackEnter = MIN2(vehLength, newPos);
        integratedLengthOnLane += (timeBeforeEnterBack - timeBeforeEnter) * (lengthOnLaneAtBackEnter + lengthOnLaneAtStepStart) * 0.5;
        integratedLengthOnLane += (timeBeforeLeaveFront - timeBeforeEnterBack) * vehLength;
        integratedLengthOnLane += (timeBeforeLeave - timeBeforeLeaveFront) * (vehLength + lengthOnLaneAtStepEnd) * 0.5;
    } else {
        double lengthOnLaneAtLeaveFront;
        if (timeBeforeLeaveFront == timeBeforeEnter) {
            lengthOnLaneAtLeaveFront = lengthOnLaneAtStepStart;
        } else if (timeBeforeLeaveFront == timeBeforeLeave) {
            lengthOnLaneAtLeaveFront = lengthOnLaneAtStepEnd;
        } else {
            lengthOnLaneAtLeaveFront = myLaneLength;
        }
        integratedLengthOnLane += (timeBeforeLeaveFront - timeBeforeEnter) * (lengthOnLaneAtLeaveFront + lengthOnLaneAtStepStart) * 0.5;
        integratedLengthOnLane += (timeBeforeEnterBack - timeBeforeLeaveFront) * lengthOnLaneAtLeaveFront;
        integratedLengthOnLane += (timeBeforeLeave - timeBeforeEnterBack) * (lengthOnLaneAtLeaveFront + lengthOnLaneAtStepEnd) * 0.5;
    }
    const double meanLengthOnLane = integratedLengthOnLane / TS;
    const double travelledDistanceFrontOnLane = frontOnLane * newSpeed;
    const double travelledDistanceVehicleOnLane = timeOnLane * newSpeed;
    const double meanSpeedVehicleOnLane = (enterSpeed + leaveSpeed) / 2.;
    // ---- MSLaneMeanDataValues::notifyMoveInternal
    md_add(c, lane, MD_SAMPLE, timeOnLane);
    md_add(c, lane, MD_DIST, travelledDistanceVehicleOnLane);
    md_add(c, lane, MD_LENGTH_SUM, vehLength * timeOnLane);
    md_add(c, lane, MD_OCCUPATION, meanLengthOnLane * TS);
    if (meanSpeedVehicleOnLane < c.md_halt_speed) {
        md_add(c, lane, MD_WAIT, timeOnLane);
    }
    if (vmax > 0) {
        md_add(c, lane, MD_TIMELOSS, timeOnLane * MAX2(0.0, vmax - meanSpeedVehicleOnLane) / vmax);
    }
    md_add(c, lane, MD_FRONT_SAMPLE, frontOnLane);
    md_add(c, lane, MD_FRONT_DIST, travelledDistanceFrontOnLane);
    atomic_min_double(&c.md_minlen[lane], veh_length_with_gap(c, v));
}

}  // namespace sb2
