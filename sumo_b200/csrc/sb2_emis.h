/* sb2_emis.h -- the HBEFA emission device: MSDevice_Emissions::notifyMove (src/microsim/devices/MSDevice_Emissions.cpp:122-126)
 * = myEmissions.addScaled(PollutantsInterface::computeAll(class, newSpeed, acceleration, slope, params), TS)
 * with computeAll = 7 x Helper::compute (utils/emissions/PollutantsInterface.cpp:361-366): HelpersHBEFA3::compute
 * (HelpersHBEFA3.h:97-116), HelpersHBEFA4::compute (HelpersHBEFA4.h:96-121), coasting cut-off
 * PollutantsInterface::Helper::getCoastingDecel (PollutantsInterface.cpp:197-210).
 * Pollutant order = PollutantsInterface::EmissionType: CO2, CO, HC, FUEL, NO_X, PM_X, ELEC.
 * The coefficient rows come from the reference's own tables through the flatten shim (vt_em_coef [nT][7][7]); flat
 * networks only (slope 0: sin(DEG2RAD(0)) = 0 exactly), engine never off (no stops).  All arithmetic is + - * / in the
 * reference's evaluation order: bit-exact.
 */
#pragma once
#include "sb2_types.h"

namespace sb2 {

enum { EM_NONE = 0, EM_HBEFA3 = 3, EM_HBEFA4 = 4, EM_N = 7, EM_ELEC = 6 };

// PollutantsInterface::Helper::getCoastingDecel for slope 0
SB2_HD double em_coastingDecel(double mass, double incl, double v) {
    const double SPEED_DCEL_MIN = 10 / 3.6;   // PHEMlightdllV5::Constants::SPEED_DCEL_MIN (foreign/PHEMlight/V5/cpp/Constants.cpp:33)
    if (v < SPEED_DCEL_MIN) {
        return v / SPEED_DCEL_MIN * MIN2(0., incl * SPEED_DCEL_MIN + 0.00001066 * mass + -0.38347107 - 20.0 * incl - 0.);
    }
    return MIN2(0., incl * v + 0.00001066 * mass + -0.38347107 - 20.0 * incl - 0.);
}

// one notifyMove: adds TS * rate to the 7 accumulators of vehicle v (pollutant-major: veh_em[e * nV + v])
SB2_HD void veh_emissions_notifyMove(const Ctx& c, int v, double newSpeed, double a) {
    if (!c.veh_emissions[v]) {
        return;
    }
    const int vt = c.veh_type[v];
    const int model = c.vt_em_model[vt];
    const double* coef = c.vt_em_coef + (size_t)vt * EM_N * 7;
    const bool coasting = newSpeed > .5 /* PollutantsInterface::Helper::ZERO_SPEED_ACCURACY, PollutantsInterface.cpp:45 */ && a < em_coastingDecel(c.vt_em_mass[vt], c.vt_em_incl[vt], newSpeed);
    const double a2 = a + 9.80665 * 0.;   // GRAVITY * sin(DEG2RAD(slope)), slope 0
    const double vv = newSpeed;
    for (int e = 0; e < EM_N; e++) {
        const double* f = coef + e * 7;
        double r;
        if (coasting) {
            r = 0.;
        } else if (model == EM_HBEFA4) {
            r = (f[0] + f[1] * vv + f[2] * a2 + f[3] * vv * vv + f[4] * vv * vv * vv + f[5] * a2 * vv + f[6] * a2 * vv * vv) / c.vt_em_scale[(size_t)vt * EM_N + e];
            if (e != EM_ELEC) {
                r = MAX2(0.0, r);
            }
        } else {   // EM_HBEFA3
            r = e == EM_ELEC ? 0. : MAX2((f[0] + f[1] * a2 * vv + f[2] * a2 * a2 * vv + f[3] * vv + f[4] * vv * vv + f[5] * vv * vv * vv) / c.vt_em_scale[(size_t)vt * EM_N + e], 0.);
        }
        c.veh_em[(size_t)e * c.nV + v] += c.TS * r;
    }
}

}  // namespace sb2
