/* sb2_e1_host.h -- host side of the E1 detectors: the reduction of an interval's event list (engine.cu, tests/host_emul).
 * Kept apart from sb2_e1.h because kernels.cu includes the device headers inside a namespace (no standard headers there). */
#pragma once
#include <algorithm>
#include <vector>
#include "sb2_e1.h"

namespace sb2 {

// MSInductLoop::writeXMLOutput (MSInductLoop.cpp:456-493) for all detectors: events = the interval's passages and the pseudo events of
// the vehicles still on a detector; they are added per detector in the order (leave time, vehicle) -- host code
struct E1Result {   // == sumob200_e1_interval
    int32_t nVehContrib, nVehEntered;
    double flow, occupancy, speed, harmonicMeanSpeed, length;
};
inline void e1_sort(std::vector<E1Event>& ev) {
    std::sort(ev.begin(), ev.end(), [](const E1Event& a, const E1Event& b) {
        if (a.det != b.det) {
            return a.det < b.det;
        }
        const bool sa = a.leave == E1_STILL_ON, sb = b.leave == E1_STILL_ON;
        if (sa != sb) {
            return sb;
        }
        if (a.leave != b.leave) {
            return a.leave < b.leave;
        }
        return a.veh < b.veh;
    });
}

inline void e1_reduce(std::vector<E1Event>& ev, const int32_t* entered, int nD, const double* e1_pos, const double* e1_end,
                      SUMOTime startTime, SUMOTime stopTime, E1Result* out) {
    e1_sort(ev);
    const double t = (double)(stopTime - startTime) / 1000.;
    const double start = (double)startTime / 1000., stop = (double)stopTime / 1000.;
    size_t at = 0;
    for (int d = 0; d < nD; d++) {
        double occupancy = 0., speedSum = 0., lengthSum = 0., inverseSpeedSum = 0.;
        int contrib = 0;
        const double detLength = e1_end[d] - e1_pos[d];
        size_t i = at;
        for (; i < ev.size() && ev[i].det == d && ev[i].leave != E1_STILL_ON; i++) {
            const E1Event& x = ev[i];
            const double timeOnDetDuringInterval = x.leave - MAX2(start, x.entry);
            occupancy += MIN2(timeOnDetDuringInterval, t);
            if (!x.leftEarly) {
                const double speed = (x.length + detLength) / MAX2(x.leave - x.entry, SB2_NUMERICAL_EPS);
                speedSum += speed;
                inverseSpeedSum += 1. / speed;
                lengthSum += x.length;
                contrib++;
            }
        }
        const double flow = (double)contrib / t * 3600.;
        for (; i < ev.size() && ev[i].det == d; i++) {
            occupancy += stop - MAX2(start, ev[i].entry);
        }
        at = i;
        occupancy *= 100. / t;
        out[d].nVehContrib = contrib;
        out[d].nVehEntered = entered[d];
        out[d].flow = flow;
        out[d].occupancy = occupancy;
        out[d].speed = contrib != 0 ? speedSum / (double)contrib : -1;
        out[d].harmonicMeanSpeed = contrib != 0 ? (double)contrib / inverseSpeedSum : -1;
        out[d].length = contrib != 0 ? lengthSum / (double)contrib : -1;
    }
}

}  // namespace sb2
