/* sb2_halo.h -- edge-partitioned multi-GPU: the per-step halo exchange.
 *
 * One connected network is cut into strips of junction columns, one strip per GPU (rank).  A junction, its internal lanes and
 * the normal lanes that END at it form a block; a strip owns whole blocks ("junction-closed"), mirroring the per-lane tasking
 * of MSEdgeControl::planMovements / executeMovements (src/microsim/MSEdgeControl.cpp:150-167, 205-219).  Every rank keeps the whole
 * network and full-size vehicle arrays (same slot numbers everywhere) but only touches what Ctx::lane_vis marks:
 *
 *   AUTH   the strip itself: simulated, authoritative
 *   GHOST  the first column of each neighbour: simulated redundantly, so a vehicle crosses the cut inside ONE step through the
 *          ordinary hand-off (remove / integrate, sorted by the key of MSLane.cpp:3540-3548) and everything the strip's border
 *          vehicles read during the step (leaders on the next lane, foes and partial occupators on the neighbour junction,
 *          followers upstream of an insertion) is there; inputs are exact at the start of the step, so is the result
 *   HALO   the neighbour's second column: only read (it is what the ghost column looks at)
 *
 * At the END of every step each rank sends the final state of its own first TWO columns on either side to that neighbour
 * (its GHOST and HALO), and overwrites its ghost / halo columns with what it receives: one exchange per step, no collective
 * inside the step.  The look-ahead zone of MSVehicle::planMoveInternal (src/microsim/MSVehicle.cpp:2334, 3076) must fit in one
 * block, i.e. edges longer than maxSpeed * TS + brakeGap(maxSpeed): checked while planning (ERR_PARTITION_REACH).
 *
 * Message = header, one HaloLane per exported lane (static list), HaloVeh records (every vehicle in those lanes' queues, the
 * partial occupators on them, and teleporting vehicles whose virtual position has moved into the receiver's strip).
 * The functions below are SB2_HD: kernels in kernels.cu, plain loops in tests/host_emul (which steps several partitions in one
 * process and must reproduce the single-partition run bit for bit: tests/test_partition.py).
 */
#pragma once
#include "sb2_types.h"
#include "sb2_veh.h"
#include "sb2_e2.h"

namespace sb2 {

struct HaloHeader {
    int32_t nLanes;      // lane records that follow (= length of the export list)
    int32_t nVeh;        // vehicle records
    int32_t overflow;    // more vehicle records were needed than the message holds
    int32_t pad;
};

struct HaloLane {
    int32_t lane, first, last, nveh, npartial, partial[MAX_PARTIAL], dq_head;
    uint8_t changed, delayed, pad[2];
    double brutto;
    int64_t actkey;
};

enum { HV_QUEUE = 0, HV_PARTIAL = 1, HV_TRANSFER = 2 };

struct HaloVeh {
    int32_t slot, kind, lane, routepos, ahead, behind, nfurther, further[MAX_FURTHER], trip_waitcount, arrival_lane;
    uint8_t havetowait, moved, trip_amwaiting, pad;
    double pos, speed, prevspeed, accel, backpos, lastdist, odometer, timeloss, stopspeed, trip_routelen, trip_departpos, trip_departspeed, traci_speed;
    int64_t waiting, since_startup, jentry, jentry_ny, jconflict, depart_time, trip_wait, arrival_time, tp_proceed;
    double em[7];
};

// static description of one side's exchange, in device memory
struct HaloSide {
    const int32_t* exp_lanes; int nExp;     // lanes whose state this rank sends (its own two border columns)
    const int32_t* imp_lanes; int nImp;     // lanes whose state it receives (the neighbour's two border columns)
    int32_t* cursor;                        // [1] vehicle records reserved so far
    int32_t* slots;                         // [cap] slot | kind << 28 of every vehicle record
    int cap;
};

SB2_HD size_t halo_message_bytes(int nExp, int cap) { return sizeof(HaloHeader) + sizeof(HaloLane) * (size_t)nExp + sizeof(HaloVeh) * (size_t)cap; }
SB2_HD HaloLane* halo_lanes(void* msg) { return reinterpret_cast<HaloLane*>(static_cast<char*>(msg) + sizeof(HaloHeader)); }
SB2_HD HaloVeh* halo_vehs(void* msg, int nLanes) { return reinterpret_cast<HaloVeh*>(static_cast<char*>(msg) + sizeof(HaloHeader) + sizeof(HaloLane) * (size_t)nLanes); }

SB2_HD int halo_reserve(const HaloSide& h, int n) {
#if defined(__CUDA_ARCH__)
    return atomicAdd(h.cursor, n);
#else
    const int at = *h.cursor;
    *h.cursor += n;
    return at;
#endif
}

// ---- pack, pass 1: one exported lane -> its lane record + the slots of its vehicles
SB2_HDN void halo_pack_lane(const Ctx& c, const HaloSide& h, void* msg, int i) {
    const int l = h.exp_lanes[i];
    HaloLane& r = halo_lanes(msg)[i];
    r.lane = l;
    r.first = c.lane_first[l]; r.last = c.lane_last[l]; r.nveh = c.lane_nveh[l]; r.npartial = c.lane_npartial[l];
    for (int k = 0; k < MAX_PARTIAL; k++) {
        r.partial[k] = c.lane_partial[(size_t)l * MAX_PARTIAL + k];
    }
    r.dq_head = c.dq_head[l];
    r.changed = c.lane_changed[l]; r.delayed = c.edge_delayed[c.lane_edge[l]]; r.pad[0] = r.pad[1] = 0;
    r.brutto = c.lane_brutto[l];
    r.actkey = c.lane_actkey[l];
    const int n = r.nveh + r.npartial;
    if (n == 0) {
        return;
    }
    int at = halo_reserve(h, n);
    for (int v = c.lane_first[l]; v >= 0; v = c.veh_behind[v]) {
        if (at < h.cap) {
            h.slots[at] = v | (HV_QUEUE << 28);
        }
        at++;
    }
    for (int k = 0; k < r.npartial; k++) {
        if (at < h.cap) {
            h.slots[at] = r.partial[k] | (HV_PARTIAL << 28);
        }
        at++;
    }
}

// ---- pack, pass 1b (one thread): teleporting vehicles whose current edge lies in the receiver's strip leave this rank's list
SB2_HDN void halo_pack_transfers(const Ctx& c, const HaloSide& h, const uint8_t* imp_mask) {
    const int n = c.counters[10];
    int keep = 0;
    for (int i = 0; i < n; i++) {
        const int v = c.tp_list[i];
        const int l0 = c.edge_lane_begin[veh_succ_edge(c, v, 0)];
        if (!lane_auth(c, l0) && imp_mask[l0]) {
            const int at = halo_reserve(h, 1);
            if (at < h.cap) {
                h.slots[at] = v | (HV_TRANSFER << 28);
            }
        } else {
            c.tp_list[keep++] = v;
        }
    }
    c.counters[10] = keep;
}

// ---- pack, pass 2: vehicle record i
SB2_HDN void halo_pack_vehicle(const Ctx& c, const HaloSide& h, void* msg, int i) {
    HaloHeader* hd = static_cast<HaloHeader*>(msg);
    const int total = *h.cursor;
    if (i == 0) {
        hd->nLanes = h.nExp;
        hd->nVeh = total < h.cap ? total : h.cap;
        hd->overflow = total > h.cap;
        hd->pad = 0;
        if (total > h.cap) {
            raise(c, ERR_HALO_OVERFLOW);
        }
    }
    if (i >= total || i >= h.cap) {
        return;
    }
    const int v = h.slots[i] & 0x0fffffff;
    HaloVeh& r = halo_vehs(msg, h.nExp)[i];
    r.slot = v; r.kind = h.slots[i] >> 28;
    r.lane = c.veh_lane[v]; r.routepos = c.veh_routepos[v]; r.ahead = c.veh_ahead[v]; r.behind = c.veh_behind[v];
    r.nfurther = c.veh_nfurther[v];
    for (int k = 0; k < MAX_FURTHER; k++) {
        r.further[k] = c.veh_further[(size_t)v * MAX_FURTHER + k];
    }
    r.trip_waitcount = c.veh_trip_waitcount[v]; r.arrival_lane = c.veh_arrival_lane[v];
    r.havetowait = c.veh_havetowait[v]; r.moved = c.veh_moved[v]; r.trip_amwaiting = c.veh_trip_amwaiting[v]; r.pad = 0;
    r.pos = c.veh_pos[v]; r.speed = c.veh_speed[v]; r.prevspeed = c.veh_prevspeed[v]; r.accel = c.veh_accel[v]; r.backpos = c.veh_backpos[v];
    r.lastdist = c.veh_lastdist[v]; r.odometer = c.veh_odometer[v]; r.timeloss = c.veh_timeloss[v]; r.stopspeed = c.veh_stopspeed[v];
    r.trip_routelen = c.veh_trip_routelen[v]; r.trip_departpos = c.veh_trip_departpos[v]; r.trip_departspeed = c.veh_trip_departspeed[v];
    r.traci_speed = c.veh_traci_speed[v];
    r.waiting = c.veh_waiting[v]; r.since_startup = c.veh_since_startup[v]; r.jentry = c.veh_jentry[v]; r.jentry_ny = c.veh_jentry_ny[v];
    r.jconflict = c.veh_jconflict[v]; r.depart_time = c.veh_depart_time[v]; r.trip_wait = c.veh_trip_wait[v]; r.arrival_time = c.veh_arrival_time[v];
    r.tp_proceed = c.veh_tp_proceed[v];
    for (int p = 0; p < 7; p++) {
        r.em[p] = c.veh_emissions[v] ? c.veh_em[(size_t)p * c.nV + v] : 0.;
    }
}

// ---- unpack, pass 1: forget what this rank believed about imported lane i (its ghost results / last step's halo): every
// vehicle of the queue becomes "not visible" (-5) unless the received records put it back; a departure the ghost column made
// but the owner did not make goes back to pending (-1)
SB2_HDN void halo_clear_lane(const Ctx& c, const HaloSide& h, int i) {
    const int l = h.imp_lanes[i];
    for (int v = c.lane_first[l]; v >= 0;) {
        const int next = c.veh_behind[v];
        if (c.veh_lane[v] == l) {
            c.veh_lane[v] = (c.veh_depart_time[v] == c.t - c.DELTA_T) ? -1 : -5;
            if (SB2_OBSERVERS && (c.nE1 > 0 || c.nE2 > 0) && lane_simulated(c, l)) {
                c.veh_sim_seen[v] = c.t;   // executed here in the step that just ended: its move reminders are up to date
                if (c.veh_lane[v] == -1) {
                    veh_e1_forget(c, v);   // (a departure of the ghost column: redone below if the owner made it too)
                    veh_e2_forget(c, v);
                }
            }
        }
        v = next;
    }
    c.lane_incoming[l] = -1;
    c.lane_arrived[l] = 0;
}

// ---- unpack, pass 2: lane record i of the received message
SB2_HDN void halo_unpack_lane(const Ctx& c, const void* msg, int i) {
    const HaloLane& r = reinterpret_cast<const HaloLane*>(static_cast<const char*>(msg) + sizeof(HaloHeader))[i];
    const int l = r.lane;
    c.lane_first[l] = r.first; c.lane_last[l] = r.last; c.lane_nveh[l] = r.nveh; c.lane_npartial[l] = r.npartial;
    for (int k = 0; k < MAX_PARTIAL; k++) {
        c.lane_partial[(size_t)l * MAX_PARTIAL + k] = r.partial[k];
    }
    c.dq_head[l] = r.dq_head;
    c.lane_changed[l] = r.changed;
    if (r.delayed) {
        c.edge_delayed[c.lane_edge[l]] = 1;
    }
    c.lane_brutto[l] = r.brutto;
    c.lane_actkey[l] = r.actkey;
}

// ---- unpack, pass 3: vehicle record i
SB2_HDN void halo_unpack_vehicle(const Ctx& c, const void* msg, int i) {
    const HaloHeader* hd = static_cast<const HaloHeader*>(msg);
    if (i >= hd->nVeh) {
        return;
    }
    if (i == 0 && hd->overflow) {
        raise(c, ERR_HALO_OVERFLOW);
    }
    const HaloVeh& r = reinterpret_cast<const HaloVeh*>(static_cast<const char*>(msg) + sizeof(HaloHeader) + sizeof(HaloLane) * (size_t)hd->nLanes)[i];
    const int v = r.slot;
    if (r.kind == HV_PARTIAL && r.lane >= 0 && lane_auth(c, r.lane)) {
        return;   // a partial occupator whose front is on this rank's own lane: this rank is the authority for it
    }
    c.veh_lane[v] = r.lane; c.veh_routepos[v] = r.routepos; c.veh_ahead[v] = r.ahead; c.veh_behind[v] = r.behind;
    c.veh_nfurther[v] = r.nfurther;
    for (int k = 0; k < MAX_FURTHER; k++) {
        c.veh_further[(size_t)v * MAX_FURTHER + k] = r.further[k];
    }
    c.veh_trip_waitcount[v] = r.trip_waitcount; c.veh_arrival_lane[v] = r.arrival_lane;
    c.veh_havetowait[v] = r.havetowait; c.veh_moved[v] = r.moved; c.veh_trip_amwaiting[v] = r.trip_amwaiting;
    c.veh_pos[v] = r.pos; c.veh_speed[v] = r.speed; c.veh_prevspeed[v] = r.prevspeed; c.veh_accel[v] = r.accel; c.veh_backpos[v] = r.backpos;
    c.veh_lastdist[v] = r.lastdist; c.veh_odometer[v] = r.odometer; c.veh_timeloss[v] = r.timeloss; c.veh_stopspeed[v] = r.stopspeed;
    c.veh_trip_routelen[v] = r.trip_routelen; c.veh_trip_departpos[v] = r.trip_departpos; c.veh_trip_departspeed[v] = r.trip_departspeed;
    c.veh_traci_speed[v] = r.traci_speed;
    c.veh_waiting[v] = r.waiting; c.veh_since_startup[v] = r.since_startup; c.veh_jentry[v] = r.jentry; c.veh_jentry_ny[v] = r.jentry_ny;
    c.veh_jconflict[v] = r.jconflict; c.veh_depart_time[v] = r.depart_time; c.veh_trip_wait[v] = r.trip_wait; c.veh_arrival_time[v] = r.arrival_time;
    c.veh_tp_proceed[v] = r.tp_proceed;
    if (c.veh_emissions[v]) {
        for (int p = 0; p < 7; p++) {
            c.veh_em[(size_t)p * c.nV + v] = r.em[p];
        }
    }
    c.veh_deferred[v] = 0;
    // move reminders (E1 / E2 detectors) are not part of the record: a vehicle on a lane this rank simulates (ghost column) keeps the
    // ones its redundant execution gave it; on a lane it only reads (halo column) it has none that concern this rank; a teleporting
    // vehicle has exactly those of the lane it virtually sits on (veh_transfer_add / transfer_checkInsertions)
    if (r.kind == HV_TRANSFER) {
        veh_e1_forget(c, v);
        veh_e2_forget(c, v);
        const int vl = c.edge_lane_begin[veh_succ_edge(c, v, 0)];
        veh_e1_add_lane(c, v, vl, true, 0., -veh_length(c, v));
        veh_e2_add_lane(c, v, vl, true, -veh_length(c, v));
    } else if (SB2_OBSERVERS && (c.nE1 > 0 || c.nE2 > 0) && (r.lane < 0 || !lane_simulated(c, r.lane) || c.veh_sim_seen[v] != c.t)) {
        // only read here, or it appears on the ghost column without having been executed here in this step (it drove in from the
        // halo column, or a teleport ended there): no reminder of it concerns this rank's detectors
        veh_e1_forget(c, v);
        veh_e2_forget(c, v);
    } else if (SB2_OBSERVERS && (c.nE1 > 0 || c.nE2 > 0) && r.kind == HV_QUEUE && r.depart_time == c.t - c.DELTA_T) {
        // departed on the ghost column in the step that just ended (here too, see halo_clear_lane): addReminder + activateReminders
        veh_e1_add_lane(c, v, r.lane, true, r.pos, r.pos - veh_length(c, v));
        veh_e2_add_lane(c, v, r.lane, true, r.pos - veh_length(c, v));
    }
    if (r.kind == HV_TRANSFER) {
#if defined(__CUDA_ARCH__)
        const int at = atomicAdd(&c.counters[10], 1);
#else
        const int at = c.counters[10]++;
#endif
        c.tp_list[at] = v;
    }
}

static_assert(sizeof(HaloHeader) == 16 && sizeof(HaloLane) == 64 && sizeof(HaloVeh) == 288, "halo record layout (sumo_b200/partition.py halo_stats)");

}  // namespace sb2
