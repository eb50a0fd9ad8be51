/* host_emul -- TEST INFRASTRUCTURE, not a product path.
 *
 * Compiles the device functions of sumo_b200/csrc (sb2_*.h, the same source nvcc turns into the sm_100a kernels of
 * libsumob200.so) for the host with g++ -ffp-contract=off and runs the step phases as plain loops.  It exists because the
 * build container has no GPU: parity of the step logic against traces of the UNMODIFIED reference (oracle/_ref/sumo_flatten
 * --trace-out) is debugged here, then confirmed on the B200 through the C ABI (tests -m gpu).  libsumob200.so never links
 * or calls this.
 *
 * usage: host_emul X.flat [--trace-out X.trace] [--steps N] [--dump-veh ID]
 */
#include <cstdio>
#include <cstdlib>
#include <map>
#include <string>
#include <vector>
#include <algorithm>

#include "../../sumo_b200/csrc/sb2_build.h"
#include "../../sumo_b200/csrc/sb2_plan2.h"
#include "../../sumo_b200/csrc/sb2_exec.h"
#include "../../sumo_b200/csrc/sb2_lane.h"
#include "../../sumo_b200/csrc/sb2_insert.h"
#include "../../sumo_b200/csrc/sb2_transfer.h"
#include "../../sumo_b200/csrc/sb2_halo.h"
#include "../../sumo_b200/csrc/sb2_fcd.h"
#include "../../sumo_b200/csrc/sb2_fcdxml.h"
#include "../../sumo_b200/csrc/sb2_summary.h"
#include "../../sumo_b200/csrc/sb2_e1_host.h"

using namespace sb2;

static int gDebugVeh = -1;
static void dumpItems(const Model& m, int v, const char* tag) {
    const Ctx& c = m.h;
    const int cur = c.cur;
    int wp = -1;
    for (int i = 0; i < c.counters[7]; i++) {
        if (c.veh_order[cur][i] == v) {
            wp = i;
        }
    }
    if (wp < 0) {
        return;
    }
    fprintf(stderr, "%s t=%lld veh=%s lane=%s pos=%.6f speed=%.6f wait=%lld ndi=%d htw=%d\n", tag, (long long)c.t, m.veh_names[v].c_str(),
            c.veh_lane[v] >= 0 ? m.lane_names[c.veh_lane[v]].c_str() : "-", c.veh_pos[v], c.veh_speed[v], (long long)c.veh_waiting[v], c.veh_ndi[cur][wp], c.veh_havetowait[v]);
    if (c.veh_lane[v] >= 0) {
        fprintf(stderr, "   queue(front->back):");
        for (int x = c.lane_first[c.veh_lane[v]]; x >= 0; x = c.veh_behind[x]) {
            fprintf(stderr, " %s@%.2f", m.veh_names[x].c_str(), c.veh_pos[x]);
        }
        fprintf(stderr, " | lenInFront=%.2f brutto=%.2f npartial=%d\n", c.veh_len_in_front[v], c.lane_brutto[c.veh_lane[v]], c.lane_npartial[c.veh_lane[v]]);
    }
    for (int i = 0; i < c.veh_ndi[cur][wp]; i++) {
        const size_t r = (size_t)i * c.nV + wp;
        const int k = c.di_link[cur][r];
        fprintf(stderr, "   item %d link=%d(%s->%s state=%c) vPass=%.6f vWait=%.6f req=%d arrT=%lld arrV=%.4f arrVB=%.4f dist=%.4f accelV=%.4f avail=%.4f hadStopped=%d\n", i, k,
                k >= 0 ? m.lane_names[c.link_from[k]].c_str() : "-", k >= 0 ? m.lane_names[c.link_via[k] >= 0 ? c.link_via[k] : c.link_to[k]].c_str() : "-",
                k >= 0 ? c.link_state[k] : '-', c.di_vpass[cur][r], c.di_vwait[cur][r], c.di_request[cur][r], (long long)c.di_arrtime[cur][r], c.di_arrspeed[cur][r],
                c.di_arrspeedbraking[cur][r], c.di_dist[cur][r], c.di_accelv[cur][r], c.di_avail[cur][r], c.di_hadstopped[cur][r]);
    }
}

static void step(Model& m) {
    Ctx& c = m.h;
    for (int tl = 0; tl < c.nTL; tl++) {
        tls_advance(c, tl);
    }
    c.cur ^= 1;   // the work list and the drive items of this step go to buffer c.cur
    int nOrder = 0;
    for (int l = 0; l < c.nL; l++) {
        lane_summary(c, l, nOrder);
        if (lane_simulated(c, l)) {
            nOrder += c.lane_nveh[l];
        }
    }
    c.counters[7] = nOrder;
    c.counters[3] = 0;
    std::fill(c.link_app_head[c.cur], c.link_app_head[c.cur] + c.nK, -1);
    for (int i = 0; i < nOrder; i++) {
        veh_planMove(c, c.veh_order[c.cur][i], i);
    }
    if (gDebugVeh >= 0 && c.veh_lane[gDebugVeh] >= 0) {
        dumpItems(m, gDebugVeh, "plan");
    }
    std::fill(c.lane_npartial, c.lane_npartial + c.nL, 0);
    c.counters[9] = 0;
    c.counters[11] = 0;
    for (int i = 0; i < nOrder; i++) {
        veh_executeMove_pass1(c, i);
    }
    for (int round = 0, left = c.counters[9]; left > 0; round++) {
        left = 0;
        for (int i = 0; i < c.counters[9]; i++) {
            left += veh_execute_deferred(c, i) ? 0 : 1;
        }
        if (round == 8 && left > 0) {
            raise(c, ERR_DEFERRED_UNRESOLVED);
            break;
        }
    }
    for (int l = 0; l < c.nL; l++) {
        if (lane_simulated(c, l)) {
            lane_handoff_remove(c, l);
        }
    }
    for (int l = 0; l < c.nL; l++) {
        if (lane_simulated(c, l)) {
            lane_handoff_integrate(c, l);
        }
    }
    for (int l = 0; l < c.nL; l++) {
        if (lane_simulated(c, l)) {
            lane_detect_collisions(c, l);
        }
    }
    for (int l = 0; l < c.nL; l++) {
        if (lane_simulated(c, l)) {
            lane_sort_partials(c, l);
        }
    }
    for (int i = 0; i < 3; i++) {
        c.fleet[i] = c.fleet_next[i];
    }
    for (int l = 0; l < c.nL; l++) {
        if (c.dq_head[l] < c.dq_begin[l + 1] && lane_simulated(c, l)) {
            lane_insertVehicles(c, l);
        }
    }
    if (c.counters[10] > 0) {
        transfer_checkInsertions(c);
    }
    for (int d = 0; d < c.nE2; d++) {
        e2_detector_update(c, d);   // MSDetectorControl::updateDetectors
    }
    c.t += c.DELTA_T;
}

// ---- several partitions of one network stepped in one process (the multi-GPU scheme of sb2_halo.h without GPUs): every
// partition is a full Model with its own lane_vis; after each step the halo messages are packed, handed to the neighbour
// and unpacked exactly like sumob200_halo_pack / _unpack do on the device.  The merged trace (every vehicle reported by the
// partition that owns its lane) must equal the single-partition trace bit for bit.
// partition file (int32 words): P, nL, capacity, then per partition: lane_vis[nL], and per side (0, 1): nExp, exp..., nImp, imp...
struct EmulPart {
    Model m;
    std::vector<uint8_t> vis;
    std::vector<int32_t> exp[2], imp[2], slots[2];
    std::vector<uint8_t> impMask[2];
    int32_t cursor[2] = {0, 0};
    HaloSide side[2];
    std::vector<char> sendBuf[2];
};

static int runPartitioned(const char* flatPath, const std::string& partFile, long maxSteps, const std::string& traceOut,
                          const std::string& e1Out, long long e1Period, const std::string& e2Out, long long e2Period) {
    FILE* pf = fopen(partFile.c_str(), "rb");
    if (pf == nullptr) {
        fprintf(stderr, "cannot read %s\n", partFile.c_str());
        return 1;
    }
    std::vector<int32_t> w;
    int32_t x;
    while (fread(&x, 4, 1, pf) == 1) {
        w.push_back(x);
    }
    fclose(pf);
    size_t at = 0;
    const int P = w[at++], nL = w[at++], cap = w[at++];
    FlatFile ff;
    ff.load(flatPath);
    std::vector<EmulPart*> parts;
    for (int p = 0; p < P; p++) {
        EmulPart* e = new EmulPart();
        e->m.build(ff);
        if (e->m.h.nL != nL) {
            fprintf(stderr, "partition file does not match the scenario\n");
            return 1;
        }
        e->vis.resize(nL);
        for (int l = 0; l < nL; l++) {
            e->vis[l] = (uint8_t)w[at++];
        }
        e->m.h.lane_vis = e->vis.data();
        for (int sd = 0; sd < 2; sd++) {
            const int ne = w[at++];
            e->exp[sd].assign(w.begin() + at, w.begin() + at + ne);
            at += ne;
            const int ni = w[at++];
            e->imp[sd].assign(w.begin() + at, w.begin() + at + ni);
            at += ni;
            e->impMask[sd].assign(nL, 0);
            for (int l : e->imp[sd]) {
                e->impMask[sd][l] = 1;
            }
            e->slots[sd].resize(cap);
            HaloSide& h = e->side[sd];
            h.exp_lanes = e->exp[sd].data(); h.nExp = ne; h.imp_lanes = e->imp[sd].data(); h.nImp = ni;
            h.cursor = &e->cursor[sd]; h.slots = e->slots[sd].data(); h.cap = cap;
            e->sendBuf[sd].resize(halo_message_bytes(ne, cap));
        }
        parts.push_back(e);
    }
    FILE* f = traceOut.empty() ? nullptr : fopen(traceOut.c_str(), "wb");
    const Ctx& c0 = parts[0]->m.h;
    // detectors: every detector is reported by the partition that owns its lane; the interval lines are those of the single run
    FILE* fe1 = e1Out.empty() ? nullptr : fopen(e1Out.c_str(), "w");
    FILE* fe2 = e2Out.empty() ? nullptr : fopen(e2Out.c_str(), "w");
    auto flushE1 = [&](SUMOTime from, SUMOTime to) {
        std::vector<E1Event> ev;
        std::vector<int32_t> entered((size_t)c0.nE1, 0);
        for (EmulPart* e : parts) {
            Ctx& c = e->m.h;
            for (int v = 0; v < c.nV; v++) {
                veh_e1_collect_on_detector(c, v);
            }
            const int n = std::min(c.counters[12], c.e1_event_cap);
            ev.insert(ev.end(), c.e1_events, c.e1_events + n);
            c.counters[12] = 0;
            for (int d = 0; d < c.nE1; d++) {
                entered[d] += c.e1_entered[d];
                c.e1_entered[d] = 0;
            }
        }
        std::vector<E1Result> res((size_t)c0.nE1);
        e1_reduce(ev, entered.data(), c0.nE1, c0.e1_pos, c0.e1_end, from, to, res.data());
        for (int d = 0; d < c0.nE1; d++) {
            fprintf(fe1, "%lld %lld %d %d %.17g %.17g %.17g %.17g %.17g %d\n", (long long)from, (long long)to, d,
                    res[d].nVehContrib, res[d].flow, res[d].occupancy, res[d].speed, res[d].harmonicMeanSpeed, res[d].length, res[d].nVehEntered);
        }
    };
    auto flushE2 = [&](SUMOTime from, SUMOTime to) {
        std::vector<E2Result> res((size_t)c0.nE2), scratch((size_t)c0.nE2);
        for (EmulPart* e : parts) {
            Ctx& c = e->m.h;
            for (int d = 0; d < c.nE2; d++) {
                e2_detector_readout(c, d, c.cur, e2_owned(c, d) ? res.data() : scratch.data());
            }
        }
        for (int d = 0; d < c0.nE2; d++) {
            const E2Result& r = res[d];
            fprintf(fe2, "%lld %lld %d %.17g %d %d %d %.17g %.17g %.17g %.17g %.17g %.17g %d %.17g %d %.17g %.17g %.17g %.17g %.17g %.17g %.17g %d %.17g %d\n",
                    (long long)from, (long long)to, d, r.sampledSeconds, r.nVehEntered, r.nVehLeft, r.nVehSeen, r.meanSpeed, r.meanTimeLoss, r.meanOccupancy,
                    r.maxOccupancy, r.meanMaxJamLengthInVehicles, r.meanMaxJamLengthInMeters, r.maxJamLengthInVehicles, r.maxJamLengthInMeters,
                    r.jamLengthInVehiclesSum, r.jamLengthInMetersSum, r.meanHaltingDuration, r.maxHaltingDuration, r.haltingDurationSum,
                    r.meanIntervalHaltingDuration, r.maxIntervalHaltingDuration, r.intervalHaltingDurationSum, r.startedHalts, r.meanVehicleNumber,
                    r.maxVehicleNumber);
        }
    };
    std::vector<int> order(c0.nV);
    for (int v = 0; v < c0.nV; v++) {
        order[v] = v;
    }
    std::sort(order.begin(), order.end(), [&](int a, int b) { return c0.veh_numid[a] < c0.veh_numid[b]; });
    long long updates = 0;
    int errs = 0;
    for (long s = 0; maxSteps < 0 || s < maxSteps; s++) {
        const SUMOTime t = parts[0]->m.h.t;
        if (parts[0]->m.end >= 0 && t >= parts[0]->m.end) {
            break;
        }
        for (EmulPart* e : parts) {
            step(e->m);
        }
        // pack
        for (EmulPart* e : parts) {
            for (int sd = 0; sd < 2; sd++) {
                HaloSide& h = e->side[sd];
                if (h.nExp == 0) {
                    continue;
                }
                *h.cursor = 0;
                for (int i = 0; i < h.nExp; i++) {
                    halo_pack_lane(e->m.h, h, e->sendBuf[sd].data(), i);
                }
                halo_pack_transfers(e->m.h, h, e->impMask[sd].data());
                for (int i = 0; i < cap; i++) {
                    halo_pack_vehicle(e->m.h, h, e->sendBuf[sd].data(), i);
                }
            }
        }
        // transport + unpack: side 0 of partition p talks to side 1 of partition p - 1
        for (int p = 0; p < P; p++) {
            EmulPart* e = parts[p];
            for (int sd = 0; sd < 2; sd++) {
                const int q = sd == 0 ? p - 1 : p + 1;
                HaloSide& h = e->side[sd];
                if (q < 0 || q >= P || h.nImp == 0) {
                    continue;
                }
                const void* msg = parts[q]->sendBuf[sd ^ 1].data();
                for (int i = 0; i < h.nImp; i++) {
                    halo_clear_lane(e->m.h, h, i);
                }
                for (int i = 0; i < h.nImp; i++) {
                    halo_unpack_lane(e->m.h, msg, i);
                }
                for (int i = 0; i < cap; i++) {
                    halo_unpack_vehicle(e->m.h, msg, i);
                }
            }
        }
        // network-wide fleet aggregates (MSVehicleControl.cpp:199-216)
        double agg[3] = {parts[0]->m.h.fleet_next[0], parts[0]->m.h.fleet_next[1], parts[0]->m.h.fleet_next[2]};
        for (EmulPart* e : parts) {
            agg[0] = MAX2(agg[0], e->m.h.fleet_next[0]);
            agg[1] = MIN2(agg[1], e->m.h.fleet_next[1]);
            agg[2] = MAX2(agg[2], e->m.h.fleet_next[2]);
        }
        for (EmulPart* e : parts) {
            for (int k = 0; k < 3; k++) {
                e->m.h.fleet_next[k] = agg[k];
            }
        }
        if (fe1 != nullptr && c0.nE1 > 0 && (t + c0.DELTA_T) % e1Period == 0) {
            flushE1(t + c0.DELTA_T - e1Period, t + c0.DELTA_T);
        }
        if (fe2 != nullptr && c0.nE2 > 0 && (t + c0.DELTA_T) % e2Period == 0) {
            flushE2(t + c0.DELTA_T - e2Period, t + c0.DELTA_T);
        }
        // merged trace
        std::vector<int> owner(c0.nV, -1);
        int n = 0;
        for (int p = 0; p < P; p++) {
            const Ctx& c = parts[p]->m.h;
            for (int v = 0; v < c.nV; v++) {
                if (c.veh_lane[v] >= 0 && c.lane_vis[c.veh_lane[v]] == LV_AUTH) {
                    if (owner[v] >= 0) {
                        fprintf(stderr, "t=%lld: vehicle %d is owned by two partitions\n", (long long)t, v);
                        errs++;
                    }
                    owner[v] = p;
                    n++;
                }
            }
            if (*c.err_flags != 0) {
                fprintf(stderr, "t=%lld: partition %d raised flags 0x%x\n", (long long)t, p, *c.err_flags);
                errs++;
                *c.err_flags = 0;
            }
        }
        updates += n;
        if (f != nullptr) {
            const int64_t hdr[2] = {(int64_t)t, (int64_t)n};
            fwrite(hdr, 8, 2, f);
            for (int v : order) {
                if (owner[v] >= 0) {
                    const Ctx& c = parts[owner[v]]->m.h;
                    const int64_t ids[2] = {(int64_t)c.veh_numid[v], (int64_t)c.veh_lane[v]};
                    const double st[2] = {c.veh_pos[v], c.veh_speed[v]};
                    fwrite(ids, 8, 2, f);
                    fwrite(st, 8, 2, f);
                }
            }
        }
    }
    if (f != nullptr) {
        fclose(f);
    }
    if (fe1 != nullptr) {
        fclose(fe1);
    }
    if (fe2 != nullptr) {
        fclose(fe2);
    }
    int inserted = 0, arrived = 0, teleports = 0;
    for (EmulPart* e : parts) {
        inserted += e->m.h.counters[2];
        arrived += e->m.h.counters[1];
        teleports += e->m.h.counters[6];
    }
    fprintf(stderr, "partitions=%d vehicle-updates=%lld inserted=%d arrived=%d teleports=%d errors=%d\n", P, updates, inserted, arrived, teleports, errs);
    return errs == 0 ? 0 : 3;
}

int main(int argc, char** argv) {
    if (argc < 2) {
        fprintf(stderr, "usage: host_emul X.flat [--trace-out X.trace] [--steps N]\n");
        return 2;
    }
    std::string traceOut, drawsIn, fcdOut, emisOut, tripOut, fcdXmlOut, summaryOut, tlsSetPhase, vehSetSpeed, e1Out, mdOut, e2Out;
    long long e1Period = 60000, mdPeriod = 60000, e2Period = 60000;
    long maxSteps = -1;
    int driveItems = 0;
    std::string debugName, partFile;
    for (int i = 2; i < argc; i++) {
        const std::string a = argv[i];
        if (a == "--partitions" && i + 1 < argc) {
            partFile = argv[++i];   // written by sumo_b200/partition.py:write_partition_file
            continue;
        }
        if (a == "--trace-out" && i + 1 < argc) {
            traceOut = argv[++i];
        } else if (a == "--fcd-xml-out" && i + 1 < argc) {
            fcdXmlOut = argv[++i];   // the <timestep> elements of the reference's --fcd-output
        } else if (a == "--tls-set-phase" && i + 1 < argc) {
            tlsSetPhase = argv[++i];   // "timeMs:tlsIndex:phase;..." applied before the step at that time (sumob200_tls_set_phase)
        } else if (a == "--veh-set-speed" && i + 1 < argc) {
            vehSetSpeed = argv[++i];   // "timeMs:slot:speed;..." applied before the step at that time if the vehicle is on the road (sumob200_vehicles_set_speed)
        } else if (a == "--md-out" && i + 1 < argc) {
            mdOut = argv[++i];   // text, one line per non-internal lane and interval: the raw members of MSLaneMeanDataValues (sumob200_meandata_interval)
        } else if (a == "--md-period" && i + 1 < argc) {
            mdPeriod = atoll(argv[++i]);   // ms
        } else if (a == "--drive-items" && i + 1 < argc) {
            driveItems = atoi(argv[++i]);   // the scenario scalar drive_item_capacity (networks of very short lanes)
        } else if (a == "--e2-out" && i + 1 < argc) {
            e2Out = argv[++i];   // text, one line per detector and interval: the attributes of the reference's E2 output (sumob200_e2_interval)
        } else if (a == "--e2-period" && i + 1 < argc) {
            e2Period = atoll(argv[++i]);   // ms
        } else if (a == "--e1-out" && i + 1 < argc) {
            e1Out = argv[++i];   // text, one line per detector and interval: the attributes of the reference's E1 output (sumob200_e1_interval)
        } else if (a == "--e1-period" && i + 1 < argc) {
            e1Period = atoll(argv[++i]);   // ms
        } else if (a == "--summary-out" && i + 1 < argc) {
            summaryOut = argv[++i];   // text, one line per step: the attributes of the reference's --summary-output
        } else if (a == "--tripinfo-out" && i + 1 < argc) {
            tripOut = argv[++i];   // text, one line per arrived vehicle, at the end of the run
        } else if (a == "--emissions-out" && i + 1 < argc) {
            emisOut = argv[++i];   // same binary layout as sumo_flatten --emissions-out
        } else if (a == "--fcd-out" && i + 1 < argc) {
            fcdOut = argv[++i];    // text: "t slot lane x y angle pos speed" with FCD's 2 decimals, vehicles in slot order
        } else if (a == "--draws" && i + 1 < argc) {
            drawsIn = argv[++i];   // per-vehicle dawdle draws recorded from the reference (sumo_flatten --draws-out)
        } else if (a == "--steps" && i + 1 < argc) {
            maxSteps = atol(argv[++i]);
        } else if (a == "--debug-veh" && i + 1 < argc) {
            debugName = argv[++i];
        }
    }
    if (!partFile.empty()) {
        try {
            return runPartitioned(argv[1], partFile, maxSteps, traceOut, e1Out, e1Period, e2Out, e2Period);
        } catch (const std::exception& e) {
            fprintf(stderr, "%s\n", e.what());
            return 1;
        }
    }
    FlatFile ff;
    Model m;
    try {
        ff.load(argv[1]);
        if (driveItems > 0) {
            ff.putScalar<int32_t>("drive_item_capacity", driveItems);
        }
        m.build(ff);
    } catch (const std::exception& e) {
        fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    Ctx& c = m.h;
    for (int v = 0; v < c.nV && !m.veh_names.empty(); v++) {
        if (m.veh_names[v] == debugName) {
            gDebugVeh = v;
        }
    }
    FILE* f = traceOut.empty() ? nullptr : fopen(traceOut.c_str(), "wb");
    std::vector<int> order(c.nV);
    for (int v = 0; v < c.nV; v++) {
        order[v] = v;
    }
    std::sort(order.begin(), order.end(), [&](int a, int b) { return c.veh_numid[a] < c.veh_numid[b]; });
    FILE* ffcdxml = fcdXmlOut.empty() ? nullptr : fopen(fcdXmlOut.c_str(), "w");
    FILE* fsum = summaryOut.empty() ? nullptr : fopen(summaryOut.c_str(), "w");
    FILE* fe1 = e1Out.empty() ? nullptr : fopen(e1Out.c_str(), "w");
    FILE* fmd = mdOut.empty() ? nullptr : fopen(mdOut.c_str(), "w");
    SUMOTime e1LastFlush = c.t;
    FILE* fe2 = e2Out.empty() ? nullptr : fopen(e2Out.c_str(), "w");
    SUMOTime e2LastFlush = c.t;
    auto flushE2 = [&](SUMOTime from, SUMOTime to) {
        std::vector<E2Result> res((size_t)c.nE2);
        for (int d = 0; d < c.nE2; d++) {
            e2_detector_readout(c, d, c.cur, res.data());   // (the emulation flips the buffers at the start of a step)
            const E2Result& r = res[d];
            fprintf(fe2, "%lld %lld %d %.17g %d %d %d %.17g %.17g %.17g %.17g %.17g %.17g %d %.17g %d %.17g %.17g %.17g %.17g %.17g %.17g %.17g %d %.17g %d\n",
                    (long long)from, (long long)to, d, r.sampledSeconds, r.nVehEntered, r.nVehLeft, r.nVehSeen, r.meanSpeed, r.meanTimeLoss, r.meanOccupancy,
                    r.maxOccupancy, r.meanMaxJamLengthInVehicles, r.meanMaxJamLengthInMeters, r.maxJamLengthInVehicles, r.maxJamLengthInMeters,
                    r.jamLengthInVehiclesSum, r.jamLengthInMetersSum, r.meanHaltingDuration, r.maxHaltingDuration, r.haltingDurationSum,
                    r.meanIntervalHaltingDuration, r.maxIntervalHaltingDuration, r.intervalHaltingDurationSum, r.startedHalts, r.meanVehicleNumber,
                    r.maxVehicleNumber);
        }
    };
    SUMOTime mdLastFlush = c.t;
    auto flushMd = [&](SUMOTime from) {
        for (int l = 0; l < c.nL; l++) {
            if (lane_internal(c, l)) {
                continue;
            }
            // begin lane sampleSeconds travelledDistance waitSeconds timeLoss frontSampleSeconds frontTravelledDistance vehLengthSum occupationSum
            // minimalVehicleLength departed arrived entered left teleported
            fprintf(fmd, "%lld %d", (long long)from, l);
            for (int k = 0; k < MD_NSUMS; k++) {
                fprintf(fmd, " %.17g", c.md_sums[(size_t)k * c.nL + l]);
                c.md_sums[(size_t)k * c.nL + l] = 0.;
            }
            fprintf(fmd, " %.17g", c.md_minlen[l] == SB2_DBL_MAX ? -1. : c.md_minlen[l]);
            c.md_minlen[l] = SB2_DBL_MAX;
            for (int k = 0; k < MD_NCOUNTS; k++) {
                fprintf(fmd, " %d", c.md_counts[(size_t)k * c.nL + l]);
                c.md_counts[(size_t)k * c.nL + l] = 0;
            }
            fprintf(fmd, "\n");
        }
    };
    auto flushE1 = [&](SUMOTime from, SUMOTime to) {
        for (int v = 0; v < c.nV; v++) {
            veh_e1_collect_on_detector(c, v);
        }
        const int n = std::min(c.counters[12], c.e1_event_cap);
        std::vector<E1Event> ev(c.e1_events, c.e1_events + n);
        std::vector<E1Result> res((size_t)c.nE1);
        e1_reduce(ev, c.e1_entered, c.nE1, c.e1_pos, c.e1_end, from, to, res.data());
        c.counters[12] = 0;
        for (int d = 0; d < c.nE1; d++) {
            c.e1_entered[d] = 0;
            // begin end detector nVehContrib flow occupancy speed harmonicMeanSpeed length nVehEntered
            fprintf(fe1, "%lld %lld %d %d %.17g %.17g %.17g %.17g %.17g %d\n", (long long)from, (long long)to, d,
                    res[d].nVehContrib, res[d].flow, res[d].occupancy, res[d].speed, res[d].harmonicMeanSpeed, res[d].length, res[d].nVehEntered);
        }
    };
    FILE* femis = emisOut.empty() ? nullptr : fopen(emisOut.c_str(), "wb");
    FILE* ffcd = fcdOut.empty() ? nullptr : fopen(fcdOut.c_str(), "w");
    FILE* fdraws = drawsIn.empty() ? nullptr : fopen(drawsIn.c_str(), "rb");
    std::vector<double> replay(c.nV, 0.);
    std::map<int64_t, int> slotOfNumId;
    for (int v = 0; v < c.nV; v++) {
        slotOfNumId[c.veh_numid[v]] = v;
    }
    std::multimap<int64_t, std::pair<int, int> > phaseEvents;
    for (size_t a = 0; a < tlsSetPhase.size();) {
        size_t b = tlsSetPhase.find(';', a);
        if (b == std::string::npos) {
            b = tlsSetPhase.size();
        }
        long long tm; int tl, ph;
        if (sscanf(tlsSetPhase.substr(a, b - a).c_str(), "%lld:%d:%d", &tm, &tl, &ph) == 3) {
            phaseEvents.insert(std::make_pair((int64_t)tm, std::make_pair(tl, ph)));
        }
        a = b + 1;
    }
    std::multimap<int64_t, std::pair<int, double> > speedEvents;
    for (size_t a = 0; a < vehSetSpeed.size();) {
        size_t b = vehSetSpeed.find(';', a);
        if (b == std::string::npos) {
            b = vehSetSpeed.size();
        }
        long long tm; int slot; double sp;
        if (sscanf(vehSetSpeed.substr(a, b - a).c_str(), "%lld:%d:%lf", &tm, &slot, &sp) == 3) {
            speedEvents.insert(std::make_pair((int64_t)tm, std::make_pair(slot, sp)));
        }
        a = b + 1;
    }
    long steps = 0;
    long long updates = 0;
    int lastErr = 0;
    while ((m.end < 0 || c.t < m.end) && (maxSteps < 0 || steps < maxSteps)) {
        const SUMOTime t = c.t;
        if (fdraws != nullptr) {
            int64_t hdr[3];
            if (fread(hdr, 8, 3, fdraws) == 3) {
                for (int64_t i = 0; i < hdr[1]; i++) {
                    int64_t id; double d;
                    if (fread(&id, 8, 1, fdraws) == 1 && fread(&d, 8, 1, fdraws) == 1) {
                        replay[slotOfNumId[id]] = d;
                    }
                }
                c.rng_replay = replay.data();
                c.rng_mode = 2;
            }
        }
        for (auto it = phaseEvents.lower_bound(t); it != phaseEvents.end() && it->first == t; ++it) {
            tls_set_phase(c, it->second.first, it->second.second);
        }
        for (auto it = speedEvents.lower_bound(t); it != speedEvents.end() && it->first == t; ++it) {
            if (c.veh_lane[it->second.first] >= 0) {
                c.traci_speed_used = 1;
                c.veh_traci_speed[it->second.first] = it->second.second >= 0. ? it->second.second : -1.;
            }
        }
        step(m);
        steps++;
        int n = 0;
        for (int v = 0; v < c.nV; v++) {
            n += c.veh_lane[v] >= 0;
        }
        updates += n;
        if (f != nullptr) {
            const int64_t hdr[2] = {(int64_t)t, (int64_t)n};
            fwrite(hdr, 8, 2, f);
            for (int v : order) {
                if (c.veh_lane[v] >= 0) {
                    const int64_t ids[2] = {(int64_t)c.veh_numid[v], (int64_t)c.veh_lane[v]};
                    const double st[2] = {c.veh_pos[v], c.veh_speed[v]};
                    fwrite(ids, 8, 2, f);
                    fwrite(st, 8, 2, f);
                }
            }
        }
        if (ffcdxml != nullptr) {
            std::vector<sumob200_fcd_row> rows;
            for (int v = 0; v < c.nV; v++) {
                if (c.veh_lane[v] >= 0) {
                    const Pos2 p = veh_position(c, v);
                    sumob200_fcd_row r;
                    r.slot = v; r.lane = c.veh_lane[v]; r.pos = c.veh_pos[v]; r.speed = c.veh_speed[v];
                    r.x = p.x; r.y = p.y; r.angle = naviDegree(veh_computeAngle(c, v, p));
                    rows.push_back(r);
                }
            }
            const FcdXmlNames nm = {&m.veh_names, &m.lane_names, &m.vt_names, c.veh_type, c.veh_idrank, fcdxml_precision(c.DELTA_T)};
            std::string xml;
            fcdxml_timestep(nm, rows.data(), (int)rows.size(), t, xml, 2);
            fwrite(xml.data(), 1, xml.size(), ffcdxml);
        }
        if (fsum != nullptr) {
            SummaryPart p;
            summary_zero(p);
            for (int v = 0; v < c.nV; v++) {
                summary_vehicle(c, v, t, p);
            }
            // time inserted running waiting ended arrived collisions teleports halting stopped meanWaitingTime meanTravelTime meanSpeed meanSpeedRelative
            fprintf(fsum, "%lld %d %d %d %d %d %d %d %d %d %.17g %.17g %.17g %.17g\n", (long long)t, p.departed, p.count + p.teleporting, p.waiting, p.ended, p.ended,
                    c.counters[4], c.counters[6], p.halting, 0, p.departed != 0 ? ((double)p.delay_ms / 1000.) / (double)p.departed : -1.,
                    p.ended != 0 ? ((double)p.travel_ms / 1000.) / (double)p.ended : -1., p.count > 0 ? p.speedSum / p.count : -1.,
                    p.count > 0 ? p.relSpeedSum / p.count : -1.);
        }
        if (fe1 != nullptr && c.nE1 > 0 && (t + c.DELTA_T) % e1Period == 0) {
            // MSDetectorControl::writeOutput(step + DELTA_T): the interval [t + DELTA_T - period, t + DELTA_T)
            flushE1(t + c.DELTA_T - e1Period, t + c.DELTA_T);
            e1LastFlush = t + c.DELTA_T;
        }
        if (fe2 != nullptr && c.nE2 > 0 && (t + c.DELTA_T) % e2Period == 0) {
            flushE2(t + c.DELTA_T - e2Period, t + c.DELTA_T);
            e2LastFlush = t + c.DELTA_T;
        }
        if (fmd != nullptr && c.md_on && (t + c.DELTA_T) % mdPeriod == 0) {
            flushMd(t + c.DELTA_T - mdPeriod);
            mdLastFlush = t + c.DELTA_T;
        }
        if (femis != nullptr) {
            int64_t ne = 0;
            for (int v : order) {
                ne += c.veh_lane[v] >= 0 && c.veh_emissions[v];
            }
            const int64_t ehdr[2] = {(int64_t)t, ne};
            fwrite(ehdr, 8, 2, femis);
            for (int v : order) {
                if (c.veh_lane[v] >= 0 && c.veh_emissions[v]) {
                    const int64_t id = c.veh_numid[v];
                    fwrite(&id, 8, 1, femis);
                    for (int p = 0; p < 7; p++) {
                        fwrite(&c.veh_em[(size_t)p * c.nV + v], 8, 1, femis);
                    }
                }
            }
        }
        if (ffcd != nullptr) {
            for (int v = 0; v < c.nV; v++) {
                if (c.veh_lane[v] >= 0) {
                    const Pos2 p = veh_position(c, v);
                    fprintf(ffcd, "%.2f %d %d %.2f %.2f %.2f %.2f %.2f\n", (double)t / 1000., v, c.veh_lane[v], p.x, p.y,
                            naviDegree(veh_computeAngle(c, v, p)), c.veh_pos[v], c.veh_speed[v]);
                }
            }
        }
        if (*c.err_flags != lastErr) {
            fprintf(stderr, "t=%lld: error flags now 0x%x\n", (long long)t, *c.err_flags);
            lastErr = *c.err_flags;
        }
        // termination like MSNet::simulationState: no vehicles left and none pending
        if (m.end < 0 && n == 0 && c.counters[2] + c.counters[15] == c.nV) {
            break;
        }
    }
    if (f != nullptr) {
        fclose(f);
    }
    if (ffcd != nullptr) {
        fclose(ffcd);
    }
    if (fe1 != nullptr) {
        if (c.nE1 > 0 && c.t > e1LastFlush) {
            flushE1(e1LastFlush, c.t);   // MSDetectorControl::close: the partial interval at the end of the simulation
        }
        fclose(fe1);
    }
    if (fe2 != nullptr) {
        if (c.nE2 > 0 && c.t > e2LastFlush) {
            flushE2(e2LastFlush, c.t);
        }
        fclose(fe2);
    }
    if (fmd != nullptr) {
        if (c.md_on && c.t > mdLastFlush) {
            flushMd(mdLastFlush);   // the partial interval MSDetectorControl::close writes at the end of the simulation
        }
        fclose(fmd);
    }
    if (fsum != nullptr) {
        fclose(fsum);
    }
    if (femis != nullptr) {
        fclose(femis);
    }
    if (ffcdxml != nullptr) {
        fclose(ffcdxml);
    }
    if (!tripOut.empty()) {
        FILE* ft = fopen(tripOut.c_str(), "w");
        for (int v = 0; v < c.nV; v++) {
            if (c.veh_lane[v] == -2) {
                // slot depart departLane departPos departSpeed departDelay arrival arrivalLane arrivalPos arrivalSpeed duration routeLength waitingTime waitingCount timeLoss speedFactor
                fprintf(ft, "%d %lld %d %.17g %.17g %lld %lld %d %.17g %.17g %lld %.17g %lld %d %lld %.17g\n", v, (long long)c.veh_depart_time[v], c.veh_departlane[v],
                        c.veh_trip_departpos[v], c.veh_trip_departspeed[v], (long long)(c.veh_depart_time[v] - c.veh_depart_param[v]), (long long)c.veh_arrival_time[v],
                        c.veh_arrival_lane[v], c.veh_arrivalpos[v], c.veh_speed[v], (long long)(c.veh_arrival_time[v] - c.veh_depart_time[v]),
                        c.veh_trip_routelen[v] + c.veh_arrivalpos[v], (long long)c.veh_trip_wait[v], c.veh_trip_waitcount[v], (long long)TIME2STEPS(c.veh_timeloss[v]),
                        c.veh_speedfactor[v]);
            }
        }
        fclose(ft);
    }
    fprintf(stderr, "steps=%ld vehicle-updates=%lld inserted=%d arrived=%d collisions=%d emergencyStops=%d err=0x%x\n", steps, updates,
            c.counters[2], c.counters[1], c.counters[4], c.counters[5], *c.err_flags);
    return 0;
}
