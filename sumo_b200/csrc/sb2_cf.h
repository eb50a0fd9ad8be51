/* sb2_cf.h -- car-following algebra (semi-implicit Euler update only; MSGlobals::gSemiImplicitEulerUpdate == true).
 *
 * Restates, expression for expression (fp64, no FMA contraction: nvcc -fmad=false / g++ -ffp-contract=off):
 *   MSCFModel            src/microsim/cfmodels/MSCFModel.cpp:87-108 (brakeGap), 112-132 (freeSpeed), 178-195 (getSecureGap),
 *                        324-348 (max/minNextSpeed), 461-479 (getMinimalArrivalTime), 578-591 (getMinimalArrivalSpeedEuler),
 *                        798-802 (estimateSpeedAfterDistance), 807-884 (maximumSafeStopSpeed[Euler]), 958-1034
 *                        (maximumSafeFollowSpeed), 1038-1086 (calculateEmergencyDeceleration)
 *   MSCFModel_Krauss     src/microsim/cfmodels/MSCFModel_Krauss.cpp:127-176 (stopSpeed, followSpeed, dawdle2)
 *   MSCFModel_IDM        src/microsim/cfmodels/MSCFModel_IDM.cpp:80-134,178-269
 */
#pragma once
#include "sb2_types.h"

namespace sb2 {

#if defined(SB2_NO_IDM)
#define SB2_IS_IDM(m) false
#define SB2_IS_ORIG1(m) false
#else
#define SB2_IS_IDM(m) ((m).model() == CF_IDM)
#define SB2_IS_ORIG1(m) ((m).model() == CF_KRAUSS_ORIG1)
#endif

// The car-following parameters of one vehicle type.  With -DSB2_CF_TABLE a VIEW of the vType table (vType index + context): every use
// is a load from a few-KB table that stays in L1 and is the same address for (most of) a warp.  Holding the 13 values in
// registers for the whole of planMove costs 26 of the kernel's 32 registers, i.e. spills to local memory at two 32 B sectors
// per thread and value (profiles/r01_k_plan_ncu.txt).  Default: the by-value copy (see DESIGN.md section 5 for the measured comparison).
#if defined(__CUDA_ARCH__) && defined(SB2_CF_TABLE)
// not CSE-able on purpose: the compiler would hoist the loads and keep (spill) the values again
__device__ __forceinline__ double cf_tab(const double* p) { double v; asm volatile("ld.global.nc.f64 %0, [%1];" : "=d"(v) : "l"(p)); return v; }
#else
SB2_HD double cf_tab(const double* p) { return *p; }
#endif
#if defined(SB2_CF_TABLE)
struct CF {
    const Ctx* c;
    int vt;
    double desSpeed;   // IDM only: veh->getLane()->getVehicleMaxSpeed(veh) of the EGO vehicle's current lane (set by the caller)
    SB2_HD double TS() const { return c->TS; }
    SB2_HD double accel() const { return cf_tab(c->vt_accel + vt); }
    SB2_HD double decel() const { return cf_tab(c->vt_decel + vt); }
    SB2_HD double emergencyDecel() const { return cf_tab(c->vt_emergencydecel + vt); }
    SB2_HD double apparentDecel() const { return cf_tab(c->vt_apparentdecel + vt); }
    SB2_HD double tau() const { return cf_tab(c->vt_tau + vt); }
    SB2_HD double maxSpeed() const { return cf_tab(c->vt_maxspeed + vt); }   // vType maxSpeed
    SB2_HD double sigma() const { return cf_tab(c->vt_sigma + vt); }
    SB2_HD int model() const { return c->vt_cfmodel[vt]; }
    SB2_HD double idmDelta() const { return cf_tab(c->vt_idm_delta + vt); }
    SB2_HD double idmStepping() const { return cf_tab(c->vt_idm_stepping + vt); }
    SB2_HD double minGap() const { return cf_tab(c->vt_mingap + vt); }
    SB2_HD double length() const { return cf_tab(c->vt_length + vt); }
    SB2_HD SUMOTime startupDelay() const { return c->vt_startupdelay[vt]; }
};

SB2_HD CF make_cf(const Ctx& c, int vt) {
    CF m;
    m.c = &c; m.vt = vt; m.desSpeed = 0.;
    return m;
}
#else
struct CF {
    double TS_, accel_, decel_, emergencyDecel_, apparentDecel_, tau_, maxSpeed_, sigma_, idmDelta_, idmStepping_, minGap_, length_;
    int model_;
    SUMOTime startupDelay_;
    double desSpeed;
    SB2_HD double TS() const { return TS_; }
    SB2_HD double accel() const { return accel_; }
    SB2_HD double decel() const { return decel_; }
    SB2_HD double emergencyDecel() const { return emergencyDecel_; }
    SB2_HD double apparentDecel() const { return apparentDecel_; }
    SB2_HD double tau() const { return tau_; }
    SB2_HD double maxSpeed() const { return maxSpeed_; }
    SB2_HD double sigma() const { return sigma_; }
    SB2_HD int model() const { return model_; }
    SB2_HD double idmDelta() const { return idmDelta_; }
    SB2_HD double idmStepping() const { return idmStepping_; }
    SB2_HD double minGap() const { return minGap_; }
    SB2_HD double length() const { return length_; }
    SB2_HD SUMOTime startupDelay() const { return startupDelay_; }
};

SB2_HD CF make_cf(const Ctx& c, int vt) {
    CF m;
    m.TS_ = c.TS;
    m.accel_ = c.vt_accel[vt]; m.decel_ = c.vt_decel[vt]; m.emergencyDecel_ = c.vt_emergencydecel[vt];
    m.apparentDecel_ = c.vt_apparentdecel[vt]; m.tau_ = c.vt_tau[vt]; m.maxSpeed_ = c.vt_maxspeed[vt]; m.sigma_ = c.vt_sigma[vt];
    m.model_ = c.vt_cfmodel[vt]; m.idmDelta_ = c.vt_idm_delta[vt]; m.idmStepping_ = c.vt_idm_stepping[vt];
    m.minGap_ = c.vt_mingap[vt]; m.length_ = c.vt_length[vt]; m.startupDelay_ = c.vt_startupdelay[vt];
    m.desSpeed = 0.;
    return m;
}
#endif

// MSVehicle::Influencer::influenceSpeed (MSVehicle.cpp:493-520) for the time line libsumo::Vehicle::setSpeed installs
// (Vehicle.cpp:1924-1938: two entries with the same speed, from now to the end of time) under the default speed mode 31
// (safe speed, max acceleration and max deceleration all respected): speed = s - (s - s) * td = s, then the three bounds
SB2_HD double traci_influenceSpeed(double advised, double vSafe, double vMin, double vMax) {
    double speed = advised;
    speed = vSafe < speed ? vSafe : speed;
    speed = vMax < speed ? vMax : speed;
    speed = vMin > speed ? vMin : speed;
    return speed;
}

// x / y for y > 0 with the common x == 0 answered directly (the quotient is x itself, sign included).  Same value as the plain
// division; on the device a zero numerator sends the fp64 division into its ~60-instruction slow path (3 such calls per vehicle and
// step in k_plan before this: standing vehicles have speed 0, profiles/r01_k_plan_ncu.txt)
SB2_HD double div_pos(double x, double y) { return x == 0. ? x : x / y; }

// MSCFModel.cpp:101-108
SB2_CORE double cfk_brakeGapEuler(double TS, double speed, double decel, double headwayTime) {
    const double speedReduction = decel * TS;
    const int steps = int(div_pos(speed, speedReduction));
    return (steps * speed - speedReduction * steps * (steps + 1) / 2) * TS + speed * headwayTime;
}
SB2_HD double cf_brakeGapEuler(const CF& m, double speed, double decel, double headwayTime) { return cfk_brakeGapEuler(m.TS(), speed, decel, headwayTime); }
SB2_HD double cf_brakeGap(const CF& m, double speed, double decel, double headwayTime) { return cf_brakeGapEuler(m, speed, decel, headwayTime); }
// MSCFModel.h: brakeGap(speed) = brakeGap(speed, myDecel, myHeadwayTime)
SB2_HD double cf_brakeGap(const CF& m, double speed) { return cf_brakeGapEuler(m, speed, m.decel(), m.tau()); }

// MSCFModel.cpp:324-326 (no accel profiles)
SB2_HD double cf_maxNextSpeed(const CF& m, double speed) { return MIN2(speed + m.accel() * m.TS(), m.maxSpeed()); }
// MSCFModel.cpp:330-333
// IDM: MSCFModel_IDM.cpp:80-88
SB2_HD double cf_minNextSpeed(const CF& m, double speed) {
    const double decel = SB2_IS_IDM(m) ? MAX2(m.decel(), MIN2(m.emergencyDecel(), 1.5)) : m.decel();
    return MAX2(speed - decel * m.TS(), 0.);
}
// MSCFModel.cpp:341-343
SB2_HD double cf_minNextSpeedEmergency(const CF& m, double speed) { return MAX2(speed - m.emergencyDecel() * m.TS(), 0.); }

// MSCFModel.cpp:112-132 (Euler branch); freeSpeed(veh, speed, seen, maxSpeed, onInsertion) at :353-360
SB2_CORE double cfk_freeSpeed(double TS, double decel, double dist, double targetSpeed, bool onInsertion) {
    if (targetSpeed < 0.) {
        return targetSpeed;
    }
    const double v = targetSpeed * TS;
    if (dist < v) {
        return targetSpeed;
    }
    const double b = decel * TS * TS;
    const double y = MAX2(0.0, ((sqrt((b + 2.0 * v) * (b + 2.0 * v) + 8.0 * b * dist) - b) * 0.5 - v) / b);
    const double yFull = floor(y);
    const double exactGap = (yFull * yFull + yFull) * 0.5 * b + yFull * v + (y > yFull ? v : 0.0);
    const double fullSpeedGain = (yFull + (onInsertion ? 1. : 0.)) * (decel * TS);
    return div_pos(div_pos(MAX2(0.0, dist - exactGap), yFull + 1), TS) + fullSpeedGain + targetSpeed;
}
SB2_HD double cf_idm_freeSpeed(const CF& m, double speed, double seen, double maxSpeed);
SB2_HD double cf_freeSpeed(const CF& m, double currentSpeed, double dist, double targetSpeed, bool onInsertion) {
    if (SB2_IS_IDM(m)) {
        return cf_idm_freeSpeed(m, currentSpeed, dist, targetSpeed);
    }
    return cfk_freeSpeed(m.TS(), m.decel(), dist, targetSpeed, onInsertion);
}

// MSCFModel.cpp:798-802
SB2_CORE double cfk_estimateSpeedAfterDistance(double maxSpeed, double dist, double v, double accel) {
    return MIN2(maxSpeed, (double)sqrt(MAX2(0., 2 * dist * accel + v * v)));
}
SB2_HD double cf_estimateSpeedAfterDistance(const CF& m, double dist, double v, double accel) { return cfk_estimateSpeedAfterDistance(m.maxSpeed(), dist, v, accel); }

// MSCFModel.cpp:461-479
SB2_CORE SUMOTime cfk_getMinimalArrivalTime(double mAccel, double mDecel, double dist, double currentSpeed, double arrivalSpeed) {
    if (dist <= 0.) {
        return 0;
    }
    arrivalSpeed = MAX2(0.0, arrivalSpeed);
    const double accel = (arrivalSpeed >= currentSpeed) ? mAccel : -mDecel;
    const double accelTime = accel == 0. ? 0. : (arrivalSpeed == currentSpeed ? 0. : (arrivalSpeed - currentSpeed) / accel);   // x - x = +0, accel > 0 then
    const double accelWay = accelTime * (arrivalSpeed + currentSpeed) * 0.5;
    if (dist >= accelWay) {
        const double nonAccelWay = dist - accelWay;
        const double nonAccelSpeed = MAX3(currentSpeed, arrivalSpeed, SB2_HALTING_SPEED);
        return TIME2STEPS(accelTime + div_pos(nonAccelWay, nonAccelSpeed));
    }
    return TIME2STEPS(-(currentSpeed - sqrt(currentSpeed * currentSpeed + 2 * accel * dist)) / accel);
}
SB2_HD SUMOTime cf_getMinimalArrivalTime(const CF& m, double dist, double currentSpeed, double arrivalSpeed) {
    return cfk_getMinimalArrivalTime(m.accel(), m.decel(), dist, currentSpeed, arrivalSpeed);
}

// MSCFModel.cpp:578-591
SB2_HD double cf_getMinimalArrivalSpeedEuler(const CF& m, double dist, double currentSpeed) {
    double arrivalSpeedBraking;
    if (dist < currentSpeed) {
        arrivalSpeedBraking = SB2_INVALID_SPEED;
    } else if (2 * (dist - currentSpeed * m.tau()) * -m.decel() + currentSpeed * currentSpeed >= 0) {
        arrivalSpeedBraking = cf_estimateSpeedAfterDistance(m, dist - currentSpeed * m.tau(), currentSpeed, -m.decel());
    } else {
        arrivalSpeedBraking = m.decel();
    }
    return arrivalSpeedBraking;
}

// MSCFModel.cpp:1038-1086
SB2_CORE double cfk_calculateEmergencyDeceleration(double emergencyDecel, double gap, double egoSpeed, double predSpeed, double predMaxDecel) {
    if (gap <= 0.) {
        return emergencyDecel;
    }
    const double predBrakeDist = 0.5 * predSpeed * predSpeed / predMaxDecel;
    const double b1 = 0.5 * egoSpeed * egoSpeed / (gap + predBrakeDist);
    if (b1 <= predMaxDecel) {
        return b1;
    }
    const double b2 = 0.5 * (egoSpeed * egoSpeed - predSpeed * predSpeed) / gap;
    return b2;
}
SB2_HD double cf_calculateEmergencyDeceleration(const CF& m, double gap, double egoSpeed, double predSpeed, double predMaxDecel) {
    return cfk_calculateEmergencyDeceleration(m.emergencyDecel(), gap, egoSpeed, predSpeed, predMaxDecel);
}

// MSCFModel.cpp:860-884
SB2_CORE double cfk_maximumSafeStopSpeedEuler(double TS, double tau, double gap, double decel, double headway) {
    const double g = gap - SB2_NUMERICAL_EPS;
    if (g < 0.) {
        return 0.;
    }
    const double b = decel * TS;
    const double t = headway >= 0 ? headway : tau;
    const double s = TS;
    const double n = floor(.5 - ((t + (sqrt(((s * s) + (4.0 * ((s * (2.0 * g / b - t)) + (t * t))))) * -0.5)) / s));
    const double h = 0.5 * n * (n - 1) * b * s + n * b * t;
    const double r = (g - h) / (n * s + t);
    const double x = n * b + r;
    return x;
}
SB2_HD double cf_maximumSafeStopSpeedEuler(const CF& m, double gap, double decel, double headway) {
    return cfk_maximumSafeStopSpeedEuler(m.TS(), m.tau(), gap, decel, headway);
}

// MSCFModel.cpp:807-856
SB2_HD double cf_maximumSafeStopSpeed(const CF& m, double gap, double decel, double currentSpeed, bool /*onInsertion*/, double headway, bool relaxEmergency) {
    double vsafe = cf_maximumSafeStopSpeedEuler(m, gap, decel, headway);
    if (relaxEmergency && m.decel() != m.emergencyDecel()) {
        double origSafeDecel = (currentSpeed - vsafe) / m.TS();
        if (origSafeDecel > m.decel() + SB2_NUMERICAL_EPS) {
            double safeDecel = SB2_EMERGENCY_DECEL_AMPLIFIER * cf_calculateEmergencyDeceleration(m, gap, currentSpeed, 0., 1);
            safeDecel = MAX2(safeDecel, m.decel());
            safeDecel = MIN2(safeDecel, origSafeDecel);
            vsafe = currentSpeed - safeDecel * m.TS();
            vsafe = MAX2(vsafe, 0.);
        }
    }
    return vsafe;
}

// MSCFModel.cpp:958-1034 (gComputeLC false outside the lane changer)
SB2_HD double cf_maximumSafeFollowSpeed(const CF& m, double gap, double egoSpeed, double predSpeed, double predMaxDecel, bool onInsertion, bool computeLC) {
    const double headway = m.tau();
    double x;
    if (gap >= 0 || computeLC) {
        x = cf_maximumSafeStopSpeed(m, gap + cf_brakeGap(m, predSpeed, MAX2(m.decel(), predMaxDecel), 0), m.decel(), egoSpeed, onInsertion, headway, false);
    } else {
        x = egoSpeed - m.emergencyDecel() * m.TS();
        x = MAX2(x, 0.);
    }
    if (m.decel() != m.emergencyDecel() && !onInsertion && !computeLC) {
        double origSafeDecel = (egoSpeed - x) / m.TS();
        if (origSafeDecel > m.decel() + SB2_NUMERICAL_EPS) {
            double safeDecel = SB2_EMERGENCY_DECEL_AMPLIFIER * cf_calculateEmergencyDeceleration(m, gap, egoSpeed, predSpeed, predMaxDecel);
            safeDecel = MAX2(safeDecel, m.decel());
            safeDecel = MIN2(safeDecel, origSafeDecel);
            x = egoSpeed - safeDecel * m.TS();
            x = MAX2(x, 0.);
        }
    }
    return x;
}

// ---- IDM (MSCFModel_IDM.cpp) -------------------------------------------------------------------
// pow(x, delta): delta is 4 by default -> exact repeated squaring is NOT what glibc pow returns in general, but for
// integer delta in {1,2,4} glibc's correctly rounded pow equals the correctly rounded product only approximately; the
// IDM path is therefore tolerance-checked, not bit-exact (DESIGN.md "libm").
// pow(x, delta) of MSCFModel_IDM::_v.  glibc's pow is accurate to ~0.52 ulp but not specified bit for bit, CUDA's to 2 ulp.
// For the default delta = 4 the device computes x^4 in double-double arithmetic (error-free products through explicit
// fma) and rounds once: the correctly rounded result, which is what glibc returns except when the exact value lies within
// ~0.02 ulp of a rounding boundary.  The host build (tests/host_emul) calls glibc's pow like the reference.
SB2_HD double cf_idm_pow(double x, double delta) {
#if defined(__CUDA_ARCH__)
    if (delta == 4. && x >= 0. && x < 1e75) {
        const double h = x * x;
        const double l = __fma_rn(x, x, -h);            // x^2 = h + l exactly
        const double hh = h * h;
        const double hl = __fma_rn(h, h, -hh);          // h^2 = hh + hl exactly
        const double cross = 2. * h * l;                // 2hl (l^2 is below the last bit of the sum)
        return hh + (hl + cross);
    }
#endif
    return pow(x, delta);
}

SB2_CORE double cfk_idm_v(double TS, double accel, double decel, double tau, double minGap, double idmDelta, double idmStepping,
                          double egoSpeed, double gap2pred, double predSpeed, double desSpeed, bool respectMinGap) {
    // MSCFModel_IDM.cpp:236-269
    const double twoSqrtAccelDecel = 2 * sqrt(accel * decel);
    const int iterations = MAX2(1, int(TS / idmStepping + .5));
    const double headwayTime = tau;
    double newSpeed = egoSpeed;
    double gap = gap2pred;
    if (respectMinGap) {
        gap += minGap;
    }
    for (int i = 0; i < iterations; i++) {
        const double delta_v = newSpeed - predSpeed;
        double s = MAX2(0., newSpeed * headwayTime + newSpeed * delta_v / twoSqrtAccelDecel);
        if (respectMinGap) {
            s += minGap;
        }
        gap = MAX2(SB2_NUMERICAL_EPS, gap);
        const double acc = accel * (1. - cf_idm_pow(newSpeed / MAX2(SB2_NUMERICAL_EPS, desSpeed), idmDelta) - (s * s) / (gap * gap));
        newSpeed = MAX2(0.0, newSpeed + (acc * TS) / iterations);   // ACCEL2SPEED(acc) / myIterations
        gap -= MAX2(0., ((newSpeed - predSpeed) * TS) / iterations);   // SPEED2DIST(newSpeed - predSpeed) / myIterations
    }
    return MAX2(0., newSpeed);
}
SB2_HD double cf_idm_v(const CF& m, double egoSpeed, double gap2pred, double predSpeed, double desSpeed, bool respectMinGap) {
    return cfk_idm_v(m.TS(), m.accel(), m.decel(), m.tau(), m.minGap(), m.idmDelta(), m.idmStepping(), egoSpeed, gap2pred, predSpeed, desSpeed, respectMinGap);
}

// MSCFModel_IDM::getSecureGap (MSCFModel_IDM.cpp:228-231)
SB2_HD double cf_idm_secureGap(const CF& m, double speed, double leaderSpeed) {
    const double delta_v = speed - leaderSpeed;
    return MAX2(0.0, speed * m.tau() + speed * delta_v / (2 * sqrt(m.accel() * m.decel())));
}
// MSCFModel_IDM::freeSpeed (MSCFModel_IDM.cpp:104-122)
SB2_HD double cf_idm_freeSpeed(const CF& m, double speed, double seen, double maxSpeed) {
    if (maxSpeed < 0.) {
        return maxSpeed;
    }
    const double secGap = cf_idm_secureGap(m, maxSpeed, 0);
    double vSafe;
    if (speed <= maxSpeed) {
        vSafe = cf_idm_v(m, speed, 1e6, maxSpeed, m.desSpeed, false);
    } else {
        vSafe = cf_idm_v(m, speed, MAX2(seen, secGap), 0, m.desSpeed, false);
    }
    if (seen < secGap) {
        vSafe = MIN2(vSafe, maxSpeed);
    }
    return vSafe;
}

// MSCFModel_KraussOrig1::vsafe (MSCFModel_KraussOrig1.cpp:96-113): the SK safe speed; myTauDecel = decel * tau
SB2_HD double cf_orig1_vsafe(const CF& m, double gap, double predSpeed) {
    if (predSpeed == 0 && gap < 0.01) {
        return 0;
    } else if (predSpeed == 0 && gap <= m.decel() * m.TS()) {
        return MIN2(m.decel() * m.TS(), gap / m.TS());
    }
    const double tauDecel = m.decel() * m.tau();
    return (double)(-1. * tauDecel + sqrt(tauDecel * tauDecel + (predSpeed * predSpeed) + (2. * m.decel() * gap)));
}

// ---- model dispatch -----------------------------------------------------------------------------
// stopSpeed(veh, speed, gap, decel, usage): MSCFModel_Krauss.cpp:127-134 ; relaxEmergency = (usage != FUTURE)
// IDM: MSCFModel_IDM.cpp:178-194
SB2_HD double cf_stopSpeed(const CF& m, double speed, double gap, double decel, bool future) {
    if (SB2_IS_IDM(m)) {
        if (gap < 0.01) {
            return 0;
        }
        double result = cf_idm_v(m, speed, gap, 0, m.desSpeed, false);
        if (gap > 0 && speed < SB2_NUMERICAL_EPS && result < SB2_NUMERICAL_EPS) {
            result = cf_maximumSafeStopSpeed(m, gap, decel, speed, false, m.TS(), false);
        }
        if (gap >= 0) {
            result = MIN2(result, gap / m.TS());
        }
        return result;
    }
    if (SB2_IS_ORIG1(m)) {   // MSCFModel_KraussOrig1.cpp:70-79
        return MIN2(cf_orig1_vsafe(m, gap, 0.), cf_maxNextSpeed(m, speed));
    }
    return MIN2(cf_maximumSafeStopSpeed(m, gap, decel, speed, false, m.TS() /* actionStepLengthSecs */, !future), cf_maxNextSpeed(m, speed));
}
SB2_HD double cf_stopSpeed(const CF& m, double speed, double gap) { return cf_stopSpeed(m, speed, gap, m.decel(), false); }

// followSpeed: MSCFModel_Krauss.cpp:138-152 (Euler branch)
// IDM: MSCFModel_IDM.cpp:126-133
SB2_HD double cf_followSpeed(const CF& m, double speed, double gap, double predSpeed, double predMaxDecel, bool computeLC = false) {
    if (SB2_IS_IDM(m)) {
        return cf_idm_v(m, speed, gap, predSpeed, m.desSpeed, true);
    }
    if (SB2_IS_ORIG1(m)) {   // MSCFModel_KraussOrig1.cpp:59-66
        return MIN2(cf_orig1_vsafe(m, gap, predSpeed), cf_maxNextSpeed(m, speed));
    }
    const double vsafe = cf_maximumSafeFollowSpeed(m, gap, speed, predSpeed, predMaxDecel, false, computeLC);
    const double vmax = cf_maxNextSpeed(m, speed);
    return MIN2(vsafe, vmax);
}
// MSCFModel.cpp:364-366
// IDM: MSCFModel_IDM.cpp:136-160 (the recursion of the reference is a loop here)
SB2_HD double cf_insertionFollowSpeed(const CF& m, double speed, double gap2pred, double predSpeed, double predMaxDecel) {
    if (SB2_IS_IDM(m)) {
        for (int guard = 0; guard < 1000; guard++) {
            const double s = MAX2(0., speed * m.tau() + speed * (speed - predSpeed) / (2 * sqrt(m.accel() * m.decel())));
            if (gap2pred >= s) {
                return speed;
            }
            const double speed2 = cf_followSpeed(m, speed, gap2pred, predSpeed, predMaxDecel);
            const double speed3 = cf_followSpeed(m, speed2, gap2pred, predSpeed, predMaxDecel);
            if (speed2 - speed3 < 1 * m.TS()) {
                return speed2;
            }
            speed = speed2;
        }
        return speed;
    }
    return cf_maximumSafeFollowSpeed(m, gap2pred, speed, predSpeed, predMaxDecel, true, false);
}
// MSCFModel.cpp:375-377
// IDM: MSCFModel_IDM.cpp:164-174 (base MSCFModel::insertionStopSpeed = stopSpeed(..., FUTURE))
SB2_HD double cf_insertionStopSpeed(const CF& m, double speed, double gap) {
    double result = cf_stopSpeed(m, speed, gap, m.decel(), true);
    if (SB2_IS_IDM(m)) {
        int i = 0;
        while (result - speed < -(m.decel() * m.TS()) && ++i < 10) {
            speed = result;
            result = cf_stopSpeed(m, speed, gap, m.decel(), true);
        }
    }
    return result;
}

// MSCFModel.cpp:178-195 with gComputeLC == false
SB2_HD double cf_getSecureGap(const CF& m, double speed, double leaderSpeed, double leaderMaxDecel) {
    if (SB2_IS_IDM(m)) {
        return cf_idm_secureGap(m, speed, leaderSpeed);
    }
    const double maxDecel = MAX2(m.decel(), leaderMaxDecel);
    const double bgLeader = cf_brakeGap(m, leaderSpeed, maxDecel, 0);
    return MAX2(0.0, cf_brakeGap(m, speed, m.decel(), m.tau()) - bgLeader);
}

// MSCFModel_Krauss.cpp:155-176 ; random in [0,1)
SB2_HD double cf_dawdle2(const CF& m, double speed, double sigma, double random) {
    if (SB2_IS_IDM(m)) {
        return speed;   // MSCFModel::patchSpeedBeforeLC (MSCFModel.h) returns vMax: no dawdling, no draw
    }
    if (SB2_IS_ORIG1(m)) {   // MSCFModel_KraussOrig1::dawdle (MSCFModel_KraussOrig1.cpp:82-93)
        return MAX2(0., speed - (sigma * m.accel() * random) * m.TS());
    }
    if (speed < m.accel()) {
        speed -= (sigma * speed * random) * m.TS();
    } else {
        speed -= (sigma * m.accel() * random) * m.TS();
    }
    return MAX2(0., speed);
}

}  // namespace sb2
