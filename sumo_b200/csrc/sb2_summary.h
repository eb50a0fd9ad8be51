/* sb2_summary.h -- the per-step aggregates MSNet::writeSummaryOutput prints (src/microsim/MSNet.cpp:688-730):
 * running / halting / mean speeds over the vehicles on the road (MSVehicleControl::getHaltingVehicleNo, getVehicleMeanSpeeds,
 * MSVehicleControl.cpp:545-580), vehicles waiting for insertion (MSInsertionControl::getWaitingVehicleNo = myPendingEmits,
 * MSInsertionControl.cpp:305), total departure delay (MSVehicleControl::vehicleDeparted, MSVehicleControl.cpp:200-203) and total
 * travel time of the ended vehicles (removePending, MSVehicleControl.cpp:164).
 *
 * The two totals are summed as integer milliseconds and divided once (the reference adds STEPS2TIME terms one by one in event
 * order): identical whenever the step length in seconds is a binary fraction (1, 0.5, 0.25 s), else equal to ~1e-16 relative.
 * Speed sums are reduced in a fixed tree order (deterministic), the reference adds in id order: equal to rounding, the summary
 * file prints 2 decimals.
 */
#pragma once
#include "sb2_types.h"
#include "sb2_veh.h"

namespace sb2 {

struct SummaryPart {
    int32_t count;       // on the road (and not stopped: stops are not part of the path)
    int32_t halting;
    int32_t waiting;     // depart time reached, not inserted yet
    int32_t departed;
    int32_t ended;
    int32_t teleporting; // in MSVehicleTransfer: still running (MSVehicleControl::getRunningVehicleNo), not on the road
    int64_t delay_ms;    // sum over departed vehicles of departure - STEPFLOOR(depart)
    int64_t travel_ms;   // sum over ended vehicles of arrival - departure
    double speedSum, relSpeedSum;
};

SB2_HD void summary_zero(SummaryPart& p) {
    p.count = 0; p.halting = 0; p.waiting = 0; p.departed = 0; p.ended = 0; p.teleporting = 0;
    p.delay_ms = 0; p.travel_ms = 0; p.speedSum = 0.; p.relSpeedSum = 0.;
}

SB2_HD void summary_add(SummaryPart& a, const SummaryPart& b) {
    a.count += b.count; a.halting += b.halting; a.waiting += b.waiting; a.departed += b.departed; a.ended += b.ended; a.teleporting += b.teleporting;
    a.delay_ms += b.delay_ms; a.travel_ms += b.travel_ms; a.speedSum += b.speedSum; a.relSpeedSum += b.relSpeedSum;
}

// contribution of vehicle slot v; tStep = time of the step that was executed last (the `time` attribute of the summary line)
SB2_HD void summary_vehicle(const Ctx& c, int v, SUMOTime tStep, SummaryPart& p) {
    const int lane = c.veh_lane[v];
    if (lane == -4) {
        return;   // discarded before its departure (--max-depart-delay): not loaded any more for the reference
    }
    if (lane >= 0) {
        const double speed = c.veh_speed[v];
        p.count++;
        p.speedSum += speed;
        // veh->getEdge() is the current ROUTE edge (the normal edge left last while the vehicle is on a junction's internal lane)
        const int edge = veh_succ_edge(c, v, 0);
        const double limit = edge >= 0 ? c.lane_speed[c.edge_lane_begin[edge]] : 0.;
        p.relSpeedSum += limit > 0 ? speed / limit : 0;
        if (speed < SB2_HALTING_SPEED) {
            p.halting++;
        }
    }
    // removal still pending (teleported beyond its arrival edge in MSVehicleTransfer::checkInsertions of this very step)
    const bool pendingRemoval = lane == -2 && c.veh_moved[v] == 4 && c.veh_arrival_time[v] == tStep;
    if (lane == -3 || pendingRemoval) {
        p.teleporting++;
    }
    if (lane != -1) {
        p.departed++;
        p.delay_ms += c.veh_depart_time[v] - (c.veh_depart_param[v] / c.DELTA_T) * c.DELTA_T;
        if (lane == -2 && !pendingRemoval) {
            p.ended++;
            // MSVehicleControl::removePending adds the time of the REMOVAL (MSVehicleControl.cpp:164), one step later for veh_moved 4
            p.travel_ms += c.veh_arrival_time[v] + (c.veh_moved[v] == 4 ? c.DELTA_T : 0) - c.veh_depart_time[v];
        }
    } else if (c.veh_depart[v] <= tStep) {
        p.waiting++;
    }
}

}  // namespace sb2
