/* sb2_fcd.h -- what MSFCDExport::write reads from a vehicle (src/microsim/output/MSFCDExport.cpp:119-180):
 * the cartesian position MSVehicle::getPosition (MSVehicle.cpp:1265-1301, validatePosition :1411-1428), the heading
 * MSVehicle::computeAngle (MSVehicle.cpp:1511-1571) with MSVehicle::getBackPosition (:1575-1609), as navi degrees
 * (GeomHelper::naviDegree, utils/geom/GeomHelper.cpp:192-205).  Lane geometry = MSLane::geometryPositionAtOffset
 * (MSLane.h) over PositionVector::positionAtOffset / rotationAtOffset (utils/geom/PositionVector.cpp:250-271,320-341,373-392).
 * Non-sublane model (posLat = 0), no lane change manoeuvres, flat networks (z = 0).
 * x, y, pos and speed are bit-identical to the reference; the angle goes through atan2, which is not correctly rounded on
 * either side ("parity unpinned" at full precision): the tests compare it at FCD's 2 decimals.
 */
#pragma once
#include "sb2_types.h"
#include "sb2_veh.h"

namespace sb2 {

struct Pos2 {
    double x, y;
    bool valid;
};
SB2_HD Pos2 pos2(double x, double y) { Pos2 p; p.x = x; p.y = y; p.valid = true; return p; }
SB2_HD Pos2 pos2_invalid() { Pos2 p; p.x = 0; p.y = 0; p.valid = false; return p; }
SB2_HD bool pos2_equal(const Pos2& a, const Pos2& b) { return a.valid == b.valid && (!a.valid || (a.x == b.x && a.y == b.y)); }
SB2_HD double pos2_dist(double ax, double ay, double bx, double by) {
    return sqrt((ax - bx) * (ax - bx) + (ay - by) * (ay - by) + 0. * 0.);   // Position::distanceTo, z = 0
}

// PositionVector::positionAtOffset(p1, p2, pos, 0)
SB2_HD Pos2 seg_positionAtOffset(double ax, double ay, double bx, double by, double pos) {
    const double dist = pos2_dist(ax, ay, bx, by);
    if (pos < 0. || dist < pos) {
        return pos2_invalid();
    }
    if (pos == 0.) {
        return pos2(ax, ay);
    }
    const double f = pos / dist;
    return pos2(ax + (bx - ax) * f, ay + (by - ay) * f);
}

// MSLane::geometryPositionAtOffset(offset, 0) = myShape.positionAtOffset(offset * myLengthGeometryFactor, 0)
SB2_HD Pos2 lane_geometryPositionAtOffset(const Ctx& c, int l, double offset) {
    const double pos = offset * c.lane_geomfactor[l];
    const int b = c.lane_shape_begin[l], e = c.lane_shape_begin[l + 1];
    if (e - b == 0) {
        return pos2_invalid();
    }
    const double* xy = c.shape_xy;
    if (e - b == 1) {
        return pos2(xy[2 * b], xy[2 * b + 1]);
    }
    double seenLength = 0;
    for (int i = b; i < e - 1; i++) {
        const double nextLength = pos2_dist(xy[2 * i], xy[2 * i + 1], xy[2 * i + 2], xy[2 * i + 3]);
        if (seenLength + nextLength > pos) {
            return seg_positionAtOffset(xy[2 * i], xy[2 * i + 1], xy[2 * i + 2], xy[2 * i + 3], pos - seenLength);
        }
        seenLength += nextLength;
    }
    return pos2(xy[2 * (e - 1)], xy[2 * (e - 1) + 1]);
}

// myLane->getShape().rotationAtOffset(interpolateLanePosToGeometryPos(pos)); shapes have >= 2 points
SB2_HD double lane_rotationAtOffset(const Ctx& c, int l, double offset) {
    double pos = offset * c.lane_geomfactor[l];
    const int b = c.lane_shape_begin[l], e = c.lane_shape_begin[l + 1];
    const double* xy = c.shape_xy;
    if (pos < 0) {
        double len = 0;
        for (int i = b; i < e - 1; i++) {
            len += pos2_dist(xy[2 * i], xy[2 * i + 1], xy[2 * i + 2], xy[2 * i + 3]);
        }
        pos += len;
    }
    double seenLength = 0;
    for (int i = b; i < e - 1; i++) {
        const double nextLength = pos2_dist(xy[2 * i], xy[2 * i + 1], xy[2 * i + 2], xy[2 * i + 3]);
        if (seenLength + nextLength > pos) {
            return atan2(xy[2 * i + 3] - xy[2 * i + 1], xy[2 * i + 2] - xy[2 * i]);
        }
        seenLength += nextLength;
    }
    return atan2(xy[2 * (e - 1) + 1] - xy[2 * (e - 2) + 1], xy[2 * (e - 1)] - xy[2 * (e - 2)]);
}

// MSVehicle::getPosition(0) incl. validatePosition
SB2_HD Pos2 veh_position(const Ctx& c, int v) {
    const int lane = c.veh_lane[v];
    Pos2 result = lane_geometryPositionAtOffset(c, lane, c.veh_pos[v]);
    double offset = 0;
    double lastLength = c.veh_pos[v];
    const int nf = c.veh_nfurther[v];
    for (int i = 0; !result.valid && i < nf; i++) {
        const int further = c.veh_further[(size_t)v * MAX_FURTHER + i];
        offset += lastLength;
        result = lane_geometryPositionAtOffset(c, further, c.lane_length[further] + offset);
        lastLength = c.lane_length[further];
    }
    return result;
}

// MSVehicle::getBackPosition
SB2_HD Pos2 veh_backPosition(const Ctx& c, int v) {
    const int lane = c.veh_lane[v];
    const double len = veh_length(c, v);
    if (c.veh_pos[v] >= len) {
        return lane_geometryPositionAtOffset(c, lane, c.veh_pos[v] - len);
    }
    const int nf = c.veh_nfurther[v];
    if (nf > 0) {
        const int back = c.veh_further[(size_t)v * MAX_FURTHER + nf - 1];
        const double backPos = MAX2(0.0, veh_back_pos_on_lane(c, v, back));
        return lane_geometryPositionAtOffset(c, back, backPos);
    }
    return lane_geometryPositionAtOffset(c, lane, 0);
}

// MSVehicle::computeAngle [rad]
SB2_HD double veh_computeAngle(const Ctx& c, int v, const Pos2& p1) {
    const int lane = c.veh_lane[v];
    Pos2 p2 = veh_backPosition(c, v);
    if (!p2.valid) {
        const int nf = c.veh_nfurther[v];
        if (nf > 0) {
            p2 = lane_geometryPositionAtOffset(c, c.veh_further[(size_t)v * MAX_FURTHER + nf - 1], 0);
            if (!p2.valid) {
                p2 = lane_geometryPositionAtOffset(c, lane, 0);
            }
        } else {
            p2 = lane_geometryPositionAtOffset(c, lane, 0);
        }
    }
    return !pos2_equal(p1, p2) ? atan2(p1.y - p2.y, p1.x - p2.x) : lane_rotationAtOffset(c, lane, c.veh_pos[v]);
}

// GeomHelper::naviDegree
SB2_HD double naviDegree(double angle) {
    double degree = (3.14159265358979323846 / 2. - angle) * 180. / 3.14159265358979323846;   // RAD2DEG(M_PI / 2. - angle)
    if (isinf(degree)) {
        return 0;
    }
    while (degree >= 360.) {
        degree -= 360.;
    }
    while (degree < 0.) {
        degree += 360.;
    }
    return degree;
}

}  // namespace sb2
