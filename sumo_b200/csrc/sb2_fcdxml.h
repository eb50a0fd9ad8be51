/* sb2_fcdxml.h -- host side of the FCD output: formats the binary rows the device produced
 * (sumob200_fcd_row) into the text MSFCDExport::write emits (src/microsim/output/MSFCDExport.cpp:98-300 with the default
 * attribute mask x, y, angle, type, speed, pos, lane, slope; src/microsim/devices/MSDevice_FCD.cpp:102-122), byte for byte:
 *     <timestep time="4.00">
 *         <vehicle id="v0" x="163.92" y="201.60" angle="270.00" type="car" speed="13.89" pos="32.88" lane="C2B2_0" slope="0.00"/>
 *     </timestep>
 * Vehicles appear in std::string order of their ids (the reference iterates a std::map<std::string, SUMOVehicle*>,
 * MSVehicleControl.h) = ascending veh_idrank.  Numbers: fixed, gPrecision = 2 decimals; times: time2string
 * (utils/common/SUMOTime.cpp:91-137).  Formatting is spread over host threads: at 1M vehicles the text (120 MB per step),
 * not the GPU, bounds the output rate.  Flat networks: slope="0.00".
 */
#pragma once
#include <algorithm>
#include <cstdio>
#include <string>
#include <thread>
#include <vector>

#include "../../include/sumob200.h"

namespace sb2 {

// gPrecision: 2 unless the step length is not a multiple of 10 ms, then 3 (MSFrame.cpp:1051-1053; --precision itself is not modelled)
inline int fcdxml_precision(int64_t delta_t_ms) { return delta_t_ms % 10 != 0 ? 3 : 2; }

inline std::string fcdxml_time2string(int64_t t, int precision = 2) {
    const bool neg = t < 0;
    unsigned long long a = (unsigned long long)(neg ? -t : t);
    char buf[48];
    if (precision >= 3) {
        snprintf(buf, sizeof(buf), "%s%llu.%03llu", neg ? "-" : "", a / 1000, a % 1000);
        return buf;
    }
    a = (a + 5) / 10;   // scale = 10^(3 - gPrecision)
    snprintf(buf, sizeof(buf), "%s%llu.%02llu", neg ? "-" : "", a / 100, a % 100);
    return buf;
}

struct FcdXmlNames {
    const std::vector<std::string>* veh;     // by slot
    const std::vector<std::string>* lane;    // by lane index
    const std::vector<std::string>* vtype;   // by vType index
    const int32_t* veh_type;                 // by slot
    const int32_t* veh_idrank;               // by slot
    int precision = 2;                       // fcdxml_precision(DELTA_T)
};

// appends one <timestep> element for `rows` (any order) to `out`
inline void fcdxml_timestep(const FcdXmlNames& nm, const sumob200_fcd_row* rows, int n, int64_t time_ms, std::string& out, int max_threads = 0) {
    const int prec = nm.precision >= 3 ? 3 : 2;
    if (n <= 0) {
        out += "    <timestep time=\"" + fcdxml_time2string(time_ms, prec) + "\"/>\n";
        return;
    }
    std::vector<int> order(n);
    for (int i = 0; i < n; i++) {
        order[i] = i;
    }
    std::sort(order.begin(), order.end(), [&](int a, int b) { return nm.veh_idrank[rows[a].slot] < nm.veh_idrank[rows[b].slot]; });
    int nt = max_threads > 0 ? max_threads : (int)std::thread::hardware_concurrency();
    nt = std::max(1, std::min(nt, n / 4096 + 1));
    std::vector<std::string> parts(nt);
    auto work = [&](int k) {
        std::string& s = parts[k];
        const int b = (int)((long long)n * k / nt), e = (int)((long long)n * (k + 1) / nt);
        s.reserve((size_t)(e - b) * 140);
        char num[6][40];
        for (int i = b; i < e; i++) {
            const sumob200_fcd_row& r = rows[order[i]];
            snprintf(num[0], 40, "%.*f", prec, r.x);
            snprintf(num[1], 40, "%.*f", prec, r.y);
            snprintf(num[2], 40, "%.*f", prec, r.angle);
            snprintf(num[3], 40, "%.*f", prec, r.speed);
            snprintf(num[4], 40, "%.*f", prec, r.pos);
            s += "        <vehicle id=\"";
            s += (*nm.veh)[r.slot];
            s += "\" x=\""; s += num[0];
            s += "\" y=\""; s += num[1];
            s += "\" angle=\""; s += num[2];
            s += "\" type=\""; s += (*nm.vtype)[nm.veh_type[r.slot]];
            s += "\" speed=\""; s += num[3];
            s += "\" pos=\""; s += num[4];
            s += "\" lane=\""; s += (*nm.lane)[r.lane];
            s += prec >= 3 ? "\" slope=\"0.000\"/>\n" : "\" slope=\"0.00\"/>\n";
        }
    };
    if (nt == 1) {
        work(0);
    } else {
        std::vector<std::thread> th;
        for (int k = 0; k < nt; k++) {
            th.emplace_back(work, k);
        }
        for (std::thread& t : th) {
            t.join();
        }
    }
    out += "    <timestep time=\"" + fcdxml_time2string(time_ms, prec) + "\">\n";
    for (const std::string& p : parts) {
        out += p;
    }
    out += "    </timestep>\n";
}

}  // namespace sb2
