/* sb2_types.h -- state layout of the step engine.
 *
 * One `Ctx` of raw pointers describes the whole simulation state: the flattened network (read-only after load), the
 * per-vehicle structure-of-arrays (indexed by the vehicle's load-order index, which is stable for its whole life, so a
 * grid of one thread per vehicle touches every state array fully coalesced), the per-vehicle drive items (double
 * buffered: the items written by planMove of step t stay readable as the "approach table" while planMove of step t+1
 * writes the other buffer), and per-lane queues (doubly linked through veh_ahead/veh_behind; lanes are FIFOs in the
 * non-sublane model, so hand-off is pop-front / push-back).
 *
 * All functions over a Ctx are SB2_HD: they are compiled by nvcc into the sm_100a kernels of libsumob200.so (the
 * product) and, for debugging in a container without a GPU, by g++ into tests/host_emul (test infrastructure only;
 * libsumob200.so has no CPU path).
 *
 * Reference types mirrored here: MSVehicle::State / DriveProcessItem (MSVehicle.h:1985), MSLink::ApproachingVehicleInformation
 * (MSLink.h:118), MSLink::ConflictInfo (MSLink.h:196), MSLane vehicle containers (MSLane.h myVehicles/myPartialVehicles).
 */
#pragma once
#include <stdint.h>
#include <math.h>
#include <float.h>

#if defined(__CUDACC__)
#define SB2_HD __host__ __device__ __forceinline__
/* entry functions called once per kernel: inlined, so that the members of the __grid_constant__ Ctx are read as
 * constant-bank operands (a noinline callee would get a generic pointer to the parameter space instead) */
#define SB2_HDN __host__ __device__ __forceinline__
/* pure scalar arithmetic cores that are used from dozens of call sites (fp64 division / sqrt expand to long instruction
 * sequences).  Round 1 measured real calls as faster (k_plan was ~8 k instructions with a 1.1 KB stack); with the per-thread
 * DriveItem array gone (round 2) inlining wins again: k_plan 1.295 -> 1.248 ms (profiles/r02_occupancy_sweep.txt).
 * -DSB2_CORE_CALL brings the calls back. */
#if defined(SB2_CORE_CALL)
#define SB2_CORE __host__ __device__ __noinline__
#else
#define SB2_CORE __host__ __device__ __forceinline__
#endif
/* rarely executed, register-hungry blocks (right-of-way check against the approach tables): out of line, so that their
 * register demand does not set the allocation of the whole kernel; they get the Ctx by reference (generic loads) */
#if !defined(SB2_RARE_CALL)
#define SB2_RARE __host__ __device__ __forceinline__
#else
#define SB2_RARE __host__ __device__ __noinline__
#endif
#else
#define SB2_HD inline
#define SB2_HDN inline
#define SB2_CORE inline
#define SB2_RARE inline
#endif

namespace sb2 {

typedef int64_t SUMOTime;

// src/config.h.cmake:236-243, utils/common/StdDefs.h:52, MSVehicle.cpp:118-127, SUMOTime.h:34
#define SB2_NUMERICAL_EPS 0.001
#define SB2_POSITION_EPS 0.1
#define SB2_SPEED_EPS 0.01
#define SB2_HALTING_SPEED 0.1
#define SB2_NUMERICAL_EPS_SPEED (0.1 * SB2_NUMERICAL_EPS * c.TS)
#define SB2_CRLL_LOOK_AHEAD 5
#define SB2_JUNCTION_BLOCKAGE_TIME 5
#define SB2_DIST_TO_STOPLINE_EXPECT_PRIORITY 1.0
#define SB2_DBL_MAX 1.7976931348623157e308
#define SB2_TIME_MAX (INT64_MAX - 1000)          /* SUMOTime_MAX */
#define SB2_TIME_MIN INT64_MIN                   /* SUMOTime_MIN */
#define SB2_INVALID_TIME INT64_MIN               /* MSLink.h INVALID_TIME = SUMOTime_MIN */
#define SB2_INVALID_SPEED 299792459.0            /* MSCFModel.h:33 */
#define SB2_NO_INTERSECTION 10000.0              /* MSLink.cpp:76 */
#define SB2_EMERGENCY_DECEL_AMPLIFIER 1.2        /* MSCFModel.h:35 */

// kernel set without move reminders (-DSB2_NO_OBSERVERS, kernels.cu): E1 / E2 detectors and edgeData compiled out
#if defined(SB2_NO_OBSERVERS)
#define SB2_OBSERVERS 0
#else
#define SB2_OBSERVERS 1
#endif
enum { MAX_DRIVE_ITEMS = 8, MAX_FURTHER = 4, MAX_PARTIAL = 4, MAX_E1_REMINDERS = 4, MAX_E2_REMINDERS = 8 };
// rows of Ctx::md_sums / md_counts (sb2_meandata.h)
enum { MD_SAMPLE = 0, MD_DIST = 1, MD_WAIT = 2, MD_TIMELOSS = 3, MD_FRONT_SAMPLE = 4, MD_FRONT_DIST = 5, MD_LENGTH_SUM = 6, MD_OCCUPATION = 7, MD_NSUMS = 8 };
enum { MD_DEPARTED = 0, MD_ARRIVED = 1, MD_ENTERED = 2, MD_LEFT = 3, MD_TELEPORTED = 4, MD_NCOUNTS = 5 };

// lane_flags / edge_flags bits (include/sumob200_flatfile.h)
enum { LF_INTERNAL = 1, LF_ROUNDABOUT = 2, LF_FRINGE = 4, LF_LANECHANGER = 8 };
// link_flags bits
enum {
    KF_KEEPCLEAR = 1, KF_HASFOES = 2, KF_AMCONT = 4, KF_ENTRY = 8, KF_CONFLICT_ENTRY = 16, KF_EXIT = 32,
    KF_EXIT_AFTER_IJ = 64, KF_INTERNAL_JUNCTION = 128, KF_INDIRECT = 256, KF_PEDFOE = 512, KF_TURNAROUND = 1024, KF_AMCONT_OFF = 2048
};
// MSLink::ConflictFlag (MSLink.h:187-193)
enum { CONFLICT_DEFAULT = 0, CONFLICT_DUMMY_MERGE = 1, CONFLICT_NO_INTERSECTION = 2, CONFLICT_STOP_AT_INTERNAL_JUNCTION = 3, CONFLICT_SIBLING_CONTINUATION = 4 };
// LinkLeader flags (MSLink.h:75-83)
enum { LL_IN_THE_WAY = 1, LL_FROM_LEFT = 2, LL_SAME_SOURCE = 4, LL_SAME_TARGET = 8 };
// error / event flags raised by kernels (never silently ignored: sumob200_step returns them)
enum {
    ERR_DRIVE_ITEM_OVERFLOW = 1, ERR_FURTHER_OVERFLOW = 2, ERR_PARTIAL_OVERFLOW = 4, ERR_COLLISION = 8, ERR_TELEPORT = 16,
    ERR_EMERGENCY_STOP = 32, ERR_NO_CONNECTION = 64, ERR_UNSUPPORTED_LINKSTATE = 128, ERR_LANE_ORDER = 256, ERR_LINKLEADER_OVERFLOW = 512, ERR_DEFERRED_UNRESOLVED = 1024, ERR_PARTITION_REACH = 2048, ERR_HALO_OVERFLOW = 4096, ERR_DETECTOR_OVERFLOW = 8192
};
// car-following model ids = SUMO_TAG_CF_* values are remapped at load time
enum { CF_KRAUSS = 0, CF_KRAUSS_ORIG1 = 1, CF_IDM = 2 };

// DepartSpeedDefinition / DepartPosDefinition values (utils/vehicle/SUMOVehicleParameter.h)
enum { DSD_DEFAULT = 0, DSD_GIVEN = 1, DSD_GIVEN_VEHROUTE = 2, DSD_RANDOM = 3, DSD_MAX = 4, DSD_DESIRED = 5, DSD_LIMIT = 6, DSD_LAST = 7, DSD_AVG = 8 };
enum { DPD_DEFAULT = 0, DPD_GIVEN = 1, DPD_GIVEN_VEHROUTE = 2, DPD_RANDOM = 3, DPD_FREE = 4, DPD_BASE = 5, DPD_LAST = 6, DPD_RANDOM_FREE = 7 };
enum { DLD_DEFAULT = 0, DLD_GIVEN = 1, DLD_FIRST_ALLOWED = 7 };

// hot part of one approach record (MSLink::ApproachingVehicleInformation, MSLink.h:118-165): what MSLink::blockedAtTime needs to
// dismiss a record, in ONE 32-byte sector (the scan of a foe link's approachers is the inner loop of MSLink::opened; with the
// fields spread over separate arrays every record cost five sectors).  The remaining fields stay in the di_* arrays.
struct AppRec {
    int32_t next;            // next record of the link's list, -1 = end
    int32_t veh;             // approaching vehicle; ~vehicle (negative) if the record has no request (willPass == false)
    int64_t arrtime;         // arrivalTime
    int64_t leavetime;       // leavingTime (MSLink::getLeaveTime at registration)
    double arrspeedbraking;  // arrivalSpeedBraking
};

// one entry of MSInductLoop::myVehicleDataCont (VehicleData, MSInductLoop.cpp:561-565) -- or, with leave == E1_STILL_ON, one entry of
// myVehiclesOnDet at the time the interval was collected
struct E1Event {
    int32_t det;
    int32_t veh;         // external index (load order)
    double entry, leave; // seconds
    double length;       // vehicle length
    int32_t leftEarly;
    int32_t pad;
};
#define E1_NOT_ON (-1e300)
#define E1_STILL_ON (-1e300)

// one MSE2Collector::MoveNotificationInfo (MSE2Collector.h:164-217) of a step, chained per detector; halt / ihalt = the vehicle's entries
// of myHaltingVehicleDurations / myIntervalHaltingVehicleDurations after the step (halt == 0: not in the maps)
struct E2Note {
    int32_t next;          // next note of the same detector, -1 = end
    int32_t veh;           // external vehicle index
    double dist;           // distToDetectorEnd
    double timeOn;         // timeOnDetector
    double lenOn;          // lengthOnDetector
    double timeLoss;
    double speed;
    int64_t halt, ihalt;   // ms
};
// per-detector accumulators of MSE2Collector (MSE2Collector.h:700-790): fp64 members [E2D_N][nE2], integer members [E2I_N][nE2],
// counters the vehicles' threads bump [E2C_N][nE2]
enum { E2D_SAMPLES = 0, E2D_TIMELOSS, E2D_SPEEDSUM, E2D_OCCSUM, E2D_MAXOCC, E2D_MEANMAXJAM_M, E2D_MAXJAM_M, E2D_JAMSUM_M, E2D_N };
enum { E2I_MEANMAXJAM_V = 0, E2I_MAXJAM_V, E2I_JAMSUM_V, E2I_TIMESAMPLES, E2I_MEANVEHNUM, E2I_MAXVEHNUM, E2I_STARTEDHALTS,
       E2I_PAST_SUM, E2I_PAST_MAX, E2I_PAST_N, E2I_IPAST_SUM, E2I_IPAST_MAX, E2I_IPAST_N, E2I_N };
enum { E2C_ENTERED = 0, E2C_LEFT, E2C_SEEN, E2C_N };
// the attributes of one <interval> of an e2 output file (MSE2Collector::writeXMLOutput); = sumob200_e2_interval_t of the C ABI
struct E2Result {
    double sampledSeconds;
    int32_t nVehEntered, nVehLeft, nVehSeen, maxJamLengthInVehicles;
    double meanSpeed, meanTimeLoss, meanOccupancy, maxOccupancy, meanMaxJamLengthInVehicles, meanMaxJamLengthInMeters, maxJamLengthInMeters;
    int32_t jamLengthInVehiclesSum, startedHalts;
    double jamLengthInMetersSum;
    double meanHaltingDuration, maxHaltingDuration, haltingDurationSum;
    double meanIntervalHaltingDuration, maxIntervalHaltingDuration, intervalHaltingDurationSum;
    double meanVehicleNumber;
    int32_t maxVehicleNumber, pad;
};

struct Ctx {
    // ---- options (MSGlobals) -------------------------------------------------------------
    double TS;                 // step length [s]
    SUMOTime DELTA_T;          // step length [ms]
    SUMOTime t;                // current time step (SIMSTEP)
    SUMOTime timeToImpatience, timeToTeleport, startupWaitThreshold, ignoreJunctionBlocker;
    double tlsYellowMinDecel;
    // fleet aggregates used by MSLane::getMaximumBrakeDist, running max/min over departed vehicles (MSVehicleControl.cpp:199-216):
    // [0] maxSpeedFactor [1] minDeceleration [2] maxMinGap.  `fleet` is the snapshot read during a step, `fleet_next` is updated
    // at departure and copied over at the begin of the next step.
    double* fleet; double* fleet_next;
    int nL, nE, nK, nV, nT, nTL;

    // ---- lanes (static) ------------------------------------------------------------------
    const double* lane_length; const double* lane_speed;
    const int32_t* lane_edge; const int32_t* lane_flags; const int32_t* lane_link_begin;
    const int32_t* lane_pred_logical; const int32_t* lane_pred_normal; const int32_t* lane_entry_link;
    const int32_t* lane_inc_begin; const int32_t* inc_lane; const int32_t* inc_link; const double* inc_length;
    const int32_t* lane_to_junction; const int32_t* lane_from_junction;
    const int32_t* edge_lane_begin;
    const int32_t* lane_numid;     // MSLane::getNumericalID (order of myChangedStateLanes, MSEdgeControl.h:296)
    const int32_t* lane_idrank;    // rank of MSLane::getID in std::string order (ComparatorIdLess, MSEdgeControl.cpp:245)
    const int32_t* lane_shape_begin; const double* shape_xy; const double* lane_geomfactor;   // MSLane::getShape, getLengthGeometryFactor
    // ---- lanes (dynamic) -----------------------------------------------------------------
    int32_t* lane_first;       // front-most vehicle (largest pos) or -1        (MSLane::myVehicles.back())
    int32_t* lane_last;        // rear-most vehicle or -1                        (MSLane::myVehicles.front())
    int32_t* lane_nveh;
    double* lane_brutto;       // myBruttoVehicleLengthSum
    int32_t* lane_npartial; int32_t* lane_partial;   // [nL*MAX_PARTIAL] partial occupators, sorted by back position
    int32_t* lane_arrived;     // set by executeMove when a vehicle of this lane's queue arrived (it may sit behind the front vehicle)
    int32_t* lane_incoming;    // head of the list of vehicles that entered this lane in executeMove (myVehBuffer)
    // MSEdgeControl::myActiveLanes (a std::list whose order decides who executes first, MSEdgeControl.cpp:118-262) as one key per
    // lane: 0 = not active, else ((step * 4 + phase) << 32 | rank), negated for lanes of multi-lane edges (push_front).  Lane A
    // executes before lane B iff key(A) < key(B).  lane_changed = myChangedStateLanes (gotActive since the last patchActiveLanes).
    int64_t* lane_actkey; uint8_t* lane_changed;
    // per-step lane summary (k_lane_summary)
    int32_t* ls_lastveh;       // getLastVehicleInformation(nullptr,0): rear-most of vehicles+partials, -1 if none
    double* ls_space;          // getSpaceTillLastStanding(...)
    uint8_t* ls_found_stopped;
    // ---- links ---------------------------------------------------------------------------
    const int32_t* link_from; const int32_t* link_to; const int32_t* link_via;
    uint8_t* link_state;       // dynamic for TLS controlled links
    const uint8_t* link_dir;
    const double* link_length; const double* link_foevis; const double* link_radius;
    const int32_t* link_flags; const int32_t* link_index; const int32_t* link_tl;
    const int32_t* link_internal_before; const int32_t* link_junction; const int32_t* link_entry; const int32_t* link_exit;
    const int32_t* link_foelink_begin; const int32_t* foelink;
    const int32_t* link_foelane_begin; const int32_t* foelane;
    const double* conf_lbc; const double* conf_size; const int32_t* conf_flag; const int32_t* conf_foeidx;
    const uint64_t* link_response;                   // [4*nK]
    const int32_t* link_contpred;                    // derived: the entry link consulted by lastWasContMajor/State, or -1
    SUMOTime* link_last_change;                      // myLastStateChange
    uint8_t* link_last_green;                        // myLastGreenState
    int32_t* link_app_head[2];                       // approach table heads (record = item*nV + work-list position), per buffer
    // ---- traffic lights ------------------------------------------------------------------
    const int32_t* tl_phase_begin; const SUMOTime* phase_dur; const int32_t* phase_state_begin; const uint8_t* phase_states;
    int32_t* tl_cur_phase; SUMOTime* tl_next_switch;
    const int32_t* tl_link_begin; const int32_t* tllink_index; const int32_t* tllink_link;
    // ---- vTypes --------------------------------------------------------------------------
    const double* vt_length; const double* vt_mingap; const double* vt_maxspeed; const double* vt_desiredmaxspeed;
    const double* vt_accel; const double* vt_decel; const double* vt_emergencydecel; const double* vt_apparentdecel;
    const double* vt_tau; const double* vt_sigma; const double* vt_impatience;
    const int32_t* vt_cfmodel; const SUMOTime* vt_startupdelay;
    const double* vt_collision_mingap_factor;
    const double* vt_idm_delta; const double* vt_idm_stepping;
    // HBEFA emission model of the type (sb2_emis.h): model id, [7][7] polynomial coefficients, [7] divisors, coasting parameters
    const int32_t* vt_em_model; const double* vt_em_coef; const double* vt_em_scale; const double* vt_em_mass; const double* vt_em_incl;
    // ---- routes --------------------------------------------------------------------------
    const int32_t* route_begin; const int32_t* route_edges;
    // ---- vehicles (static) ---------------------------------------------------------------
    const int32_t* veh_type; const int32_t* veh_route; const int32_t* veh_numid; const int32_t* veh_idrank;
    const SUMOTime* veh_depart; const SUMOTime* veh_depart_param; /* as given (tripinfo departDelay) */ const double* veh_speedfactor; const double* veh_arrivalpos; const double* veh_arrivalspeed; /* arrivalSpeed if given, else < 0 */
    const int32_t* veh_departlane;                   // resolved departure lane index
    const double* veh_departpos; const int32_t* veh_departpos_proc;
    const double* veh_departspeed; const int32_t* veh_departspeed_proc;
    // ---- vehicles (dynamic) --------------------------------------------------------------
    int32_t* veh_lane;         // -1 pending, -2 arrived, -3 teleporting (in MSVehicleTransfer, off the road), else lane index
    int32_t* veh_routepos;     // myCurrEdge - myRoute->begin()
    double* veh_pos; double* veh_speed; double* veh_prevspeed; double* veh_accel; double* veh_backpos;
    double* veh_lastdist;      // myLastCoveredDist
    SUMOTime* veh_waiting; SUMOTime* veh_since_startup;
    SUMOTime* veh_jentry; SUMOTime* veh_jentry_ny; SUMOTime* veh_jconflict;   // myJunctionEntryTime, ..NeverYield, ..ConflictEntryTime
    uint8_t* veh_havetowait;   // myHaveToWaitOnNextLink
    uint8_t* veh_moved;        // set by executeMove: 1 = entered a new lane, 2 = arrived
    SUMOTime* veh_depart_time; // myDeparture
    double* veh_odometer; double* veh_timeloss; double* veh_stopspeed;
    int32_t* veh_ahead; int32_t* veh_behind;   // lane queue links (towards larger / smaller position)
    int32_t* veh_next_incoming;
    int32_t* veh_nfurther; int32_t* veh_further;     // [nV*MAX_FURTHER] lanes the back still covers, nearest first
    // MSDevice_Tripinfo (src/microsim/devices/MSDevice_Tripinfo.cpp:182-310): myWaitingTime, myWaitingCount, myAmWaiting,
    // myRouteLength (without the arrival position), myDepartPos/Speed, myArrivalTime (-1 = not arrived), arrival lane
    SUMOTime* veh_trip_wait; int32_t* veh_trip_waitcount; uint8_t* veh_trip_amwaiting; double* veh_trip_routelen;
    double* veh_trip_departpos; double* veh_trip_departspeed; SUMOTime* veh_arrival_time; int32_t* veh_arrival_lane;
    const uint8_t* veh_emissions;   // vehicle carries the emission device
    double* veh_em;            // [7*nV] MSDevice_Emissions::myEmissions, pollutant-major (CO2, CO, HC, fuel, NOx, PMx, electricity)
    double* veh_len_in_front;  // cumulatedVehLength handed to planMove (MSLane.cpp:1621)
    int32_t* veh_qlane;        // lane whose queue holds the vehicle during this step's plan/execute (stable while veh_lane is rewritten)
    // executeMove pass 2 (sb2_exec.h): vehicles whose right of way depends on the NEW speed of a foe that the reference executes
    // earlier (MSLink::blockedAtTime reads the live speed, MSLink.cpp:1050): veh_deferred 1 = waiting, list + counters[9]
    uint8_t* veh_deferred; int32_t* deferred_list;
    // re-slotting (engine.cu reslot): the slot is where a vehicle's state lies in HBM; every few steps the on-road vehicles are
    // renumbered in lane order so that neighbours on the road are neighbours in memory.  veh_ext[slot] = the vehicle's EXTERNAL index
    // (load order: what the C ABI calls a slot -- mirror, FCD rows, set_speed, replayed draws), ext2slot the inverse.
    int32_t* veh_ext; int32_t* ext2slot;
    // MSVehicleTransfer (src/microsim/MSVehicleTransfer.cpp): vehicles that were teleported (jam / collision) wait off the road on
    // route edge veh_routepos until they fit on a lane again.  tp_list[0 .. counters[10]) = their slots (appended with an atomic,
    // sorted by numerical id in checkInsertions like the reference's std::sort), veh_tp_proceed = myProceedTime (-1 = not set)
    int32_t* tp_list; SUMOTime* veh_tp_proceed;
    // ---- E1 induction loops (MSInductLoop as a move reminder, sb2_e1.h); nE1 == 0: every array has one dummy element
    int nE1, e1_event_cap;
    const int32_t* lane_e1_begin;    // [nL+1] CSR: detectors of a lane = e1_of_lane[lane_e1_begin[l] .. lane_e1_begin[l+1])
    const int32_t* e1_of_lane;       // [nE1] detector indices sorted by lane (load order within a lane)
    const int32_t* e1_lane;          // [nE1] the detector's lane
    const double* e1_pos;            // [nE1] myPosition
    const double* e1_end;            // [nE1] myEndPosition = position + length
    int32_t* veh_e1_det;             // [MAX_E1_REMINDERS][nV] the vehicle's active reminders (MSBaseVehicle::myMoveReminders): detector or -1
    double* veh_e1_off;              // [MAX_E1_REMINDERS][nV] rem->second: length of the lanes left since the detector's lane
    double* veh_e1_entry;            // [MAX_E1_REMINDERS][nV] myVehiclesOnDet[veh] (entry time in s), E1_NOT_ON if not on the detector
    SUMOTime* veh_sim_seen;          // [nV] (detectors + edge partition) time at which halo_clear_lane found the vehicle on a lane this rank simulates
    int32_t* e1_entered;             // [nE1] myEnteredVehicleNumber of the running interval
    struct E1Event* e1_events;       // [e1_event_cap] myVehicleDataCont of all detectors; length = counters[12]
    // ---- E2 lane area detectors on one lane (MSE2Collector as a move reminder, sb2_e2.h); nE2 == 0: one dummy element each
    int extrapolateDepartPos;        // --extrapolate-departpos
    SUMOTime maxDepartDelay;         // --max-depart-delay (ms), -1 = off: a vehicle that could not depart for longer is discarded (veh_lane -4)
    int max_items;                   // drive items per vehicle (MAX_DRIVE_ITEMS unless the scenario asks for more)
    int nE2, e2_note_cap;
    const int32_t* lane_e2_begin;    // [nL+1] CSR like lane_e1_begin
    const int32_t* e2_of_lane;       // [nE2]
    const int32_t* e2_lane;          // [nE2] the detector's lane
    const double* e2_start;          // [nE2] myStartPos
    const double* e2_length;         // [nE2] myDetectorLength
    const double* e2_jam_speed;      // [nE2] myJamHaltingSpeedThreshold
    const double* e2_jam_dist;       // [nE2] myJamDistanceThreshold
    const int64_t* e2_jam_time;      // [nE2] myJamHaltingTimeThreshold (ms)
    int32_t* veh_e2_det;             // [MAX_E2_REMINDERS][nV] detector of the reminder or -1 (myVehicleInfos[veh] exists)
    double* veh_e2_off;              // [MAX_E2_REMINDERS][nV] rem->second
    int32_t* veh_e2_entered;         // [MAX_E2_REMINDERS][nV] VehicleInfo::hasEntered
    struct E2Note* e2_notes[2];      // [e2_note_cap] myMoveNotifications of all detectors, of this step ([cur]) and the previous one;
    int32_t* e2_head[2];             // [nE2] first note of the detector per buffer; number of notes per buffer = counters[13 + buffer]
    double* e2_acc_d;                // [E2D_N][nE2]
    int64_t* e2_acc_i;               // [E2I_N][nE2]
    int32_t* e2_cnt;                 // [E2C_N][nE2]
    // ---- edgeData / laneData (MSMeanData_Net, sb2_meandata.h); md_on == 0: one dummy element each
    int md_on;
    double md_halt_speed;            // haltingSpeedThreshold of the <edgeData> definition (default POSITION_EPS)
    double* md_sums;                 // [MD_NSUMS][nL] sampleSeconds, travelledDistance, waitSeconds, timeLoss, frontSampleSeconds, frontTravelledDistance, vehLengthSum, occupationSum
    double* md_minlen;               // [nL] minimalVehicleLength (SB2_DBL_MAX = INVALID_DOUBLE)
    int32_t* md_counts;              // [MD_NCOUNTS][nL] nVehDeparted, nVehArrived, nVehEntered, nVehLeft, nVehTeleported
    uint8_t* edge_delayed;     // MSEdge::myAmDelayed (a vehicle has entered the edge at some time, MSEdge.h:750)
    // ---- multi-GPU edge partition (sb2_halo.h): what this rank does with each lane.  nullptr = one partition owning everything.
    //   LV_AUTH  lanes of the rank's own junction columns: simulated here, results authoritative (counted, reported, exported)
    //   LV_GHOST lanes of the neighbouring rank's first column: simulated here too (so that vehicles cross the cut inside one
    //            step without a mid-step exchange), results overwritten by the owner's at the end of the step
    //   LV_HALO  lanes of the neighbour's second column: not simulated, only read (state received at the end of every step)
    //   LV_NONE  everything further away: never touched
    const uint8_t* lane_vis;
    // the visible lanes / links of a partition form one index range (lanes are numbered junction block by junction block, column by
    // column): the per-lane kernels are launched over [lane_lo, lane_hi) only.  One partition: the whole network.
    int lane_lo, lane_hi, link_lo, link_hi;
    // lane-ordered work list of this step: the vehicles on the road, lane by lane, front -> back (rebuilt by k_pre).
    // k_plan / k_execute give thread i the vehicle veh_order[i], so the 32 threads of a warp work on the same or on
    // neighbouring lanes, see the same links and foes and take the same branches.  Two segments: [0, counters[7]) = lanes whose
    // first link is not red/yellow, [nV - counters[3], nV) = lanes facing red/yellow (filled backwards), see k_pre.
    // Double buffered like the drive items (index c.cur): a drive-item record of step t is addressed by the vehicle's
    // POSITION wp in that step's work list (record = item * nV + wp), so the plan kernel writes and the execute kernel reads the
    // records fully coalesced; veh_order[buf][wp] gives the vehicle back (approach tables of the previous step use buf = cur ^ 1).
    int32_t* veh_order[2];
    // drive items, double buffered [2][max_items*nV], item-major (item i of the vehicle at work-list position wp at i*nV+wp);
    // veh_ndi[buf][wp] likewise
    int32_t* veh_ndi[2];
    int32_t* di_link[2]; uint8_t* di_request[2]; uint8_t* di_hadstopped[2];
    double* di_vpass[2]; double* di_vwait[2]; double* di_arrspeed[2]; double* di_arrspeedbraking[2];
    double* di_dist[2]; double* di_accelv[2]; double* di_avail[2];
    SUMOTime* di_arrtime[2];
    double* di_regspeed[2];              // approach record: approaching->getSpeed() at registration
    SUMOTime* di_regwaiting[2];          // approach record: waiting time at registration
    AppRec* ap_rec[2];                   // approach record (hot part) of every drive item with a link; list head: link_app_head
    int cur;                             // buffer written by this step's planMove
    // ---- departure queues (per departure lane, in load order) ----------------------------
    const int32_t* dq_begin; const int32_t* dq_veh; int32_t* dq_head;   // [nL+1], [nV], [nL]
    // ---- random numbers for the dawdle step (MSCFModel_Krauss::dawdle2): 0 = none needed (all sigma == 0),
    // 1 = counter-based generator keyed by (seed, vehicle, step), 2 = replay of host-supplied draws (rng_replay[v])
    int rng_mode; uint64_t rng_seed; const double* rng_replay;
    // ---- TraCI speed advice (libsumo::Vehicle::setSpeed -> MSVehicle::Influencer speed time line with one constant entry):
    // veh_traci_speed[v] >= 0 = advised speed, < 0 = none.  traci_speed_used == 0 (no advice given in this run yet) lets the
    // per-vehicle kernels skip the array
    int traci_speed_used; double* veh_traci_speed;
    // ---- step results --------------------------------------------------------------------
    int32_t* err_flags;        // [1] OR of ERR_*
    int32_t* counters;         // [16]: 7 = length of the work list, 8 = halting vehicles of this step; 0 running, 1 arrived(total), 2 inserted(total), 3 halting, 4 collisions, 5 emergency stops, 6 teleports, 9 deferred vehicles of this step, 10 length of tp_list, 11 removals pending until the next step, 12 length of e1_events, 13 / 14 notes in e2_notes[0] / [1], 15 vehicles discarded before departure (total)
};

SB2_HD double MIN2(double a, double b) { return a < b ? a : b; }
SB2_HD double MAX2(double a, double b) { return a > b ? a : b; }
SB2_HD double MIN3(double a, double b, double c) { return MIN2(c, a < b ? a : b); }
SB2_HD double MAX3(double a, double b, double c) { return MAX2(c, a > b ? a : b); }
SB2_HD SUMOTime MAX2T(SUMOTime a, SUMOTime b) { return a > b ? a : b; }
SB2_HD SUMOTime MIN2T(SUMOTime a, SUMOTime b) { return a < b ? a : b; }

// SUMOTime.h:58 -- static_cast<SUMOTime> of a double truncates towards zero
SB2_HD SUMOTime TIME2STEPS(double x) { return (SUMOTime)(x * 1000. + (x >= 0 ? 0.5 : -0.5)); }
SB2_HD double STEPS2TIME(SUMOTime x) { return (double)x / 1000.; }

SB2_HD void raise(const Ctx& c, int flag) {
#if defined(__CUDA_ARCH__)
    atomicOr(c.err_flags, flag);
#else
    *c.err_flags |= flag;
#endif
}
SB2_HD int32_t atomic_exch(int32_t* p, int32_t v) {
#if defined(__CUDA_ARCH__)
    return atomicExch(p, v);
#else
    const int32_t old = *p; *p = v; return old;
#endif
}
// Load that bypasses the (non-coherent) L1: for data another thread of the SAME kernel may have written, published with a
// __threadfence().  A plain load may hit a line this SM cached before the write and return the old contents.
template<class T> SB2_HD T ld_coherent(const T* p) {
#if defined(__CUDA_ARCH__)
    return __ldcg(p);
#else
    return *p;
#endif
}
SB2_HD void atomic_inc(int32_t* p) {
#if defined(__CUDA_ARCH__)
    atomicAdd(p, 1);
#else
    (*p)++;
#endif
}

SB2_HD void atomic_max_double(double* p, double v) {
#if defined(__CUDA_ARCH__)
    unsigned long long* a = (unsigned long long*)p;
    unsigned long long old = *a, assumed;
    do {
        assumed = old;
        if (__longlong_as_double((long long)assumed) >= v) {
            break;
        }
        old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(v));
    } while (assumed != old);
#else
    if (*p < v) { *p = v; }
#endif
}
SB2_HD void atomic_min_double(double* p, double v) {
#if defined(__CUDA_ARCH__)
    unsigned long long* a = (unsigned long long*)p;
    unsigned long long old = *a, assumed;
    do {
        assumed = old;
        if (__longlong_as_double((long long)assumed) <= v) {
            break;
        }
        old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(v));
    } while (assumed != old);
#else
    if (*p > v) { *p = v; }
#endif
}

// -DSB2_CHECK: debug build that validates indices taken from dynamic tables before they are used (dev stress runs only)
#if defined(SB2_CHECK)
SB2_HD int sb2_chk(const Ctx& c, long long x, long long lo, long long hi, int site) {
    if (x < lo || x >= hi) {
        printf("SB2_CHECK site %d: %lld not in [%lld, %lld) t=%lld\n", site, x, lo, hi, (long long)c.t);
        raise(c, 1 << 30);
        return (int)lo < 0 ? -1 : 0;
    }
    return (int)x;
}
#define SB2_CHK(x, lo, hi, site) sb2_chk(c, (x), (lo), (hi), (site))
#else
#define SB2_CHK(x, lo, hi, site) (x)
#endif

enum { LV_NONE = 0, LV_HALO = 1, LV_GHOST = 2, LV_AUTH = 3 };
SB2_HD int lane_vis(const Ctx& c, int l) { return c.lane_vis == nullptr ? LV_AUTH : c.lane_vis[l]; }
SB2_HD bool lane_auth(const Ctx& c, int l) { return c.lane_vis == nullptr || c.lane_vis[l] == LV_AUTH; }
SB2_HD bool lane_simulated(const Ctx& c, int l) { return c.lane_vis == nullptr || c.lane_vis[l] >= LV_GHOST; }
// event counters count what happens on the rank's own lanes only (the ghost column repeats the neighbour's work)
SB2_HD void count_event(const Ctx& c, int lane, int counter) {
    if (lane_auth(c, lane)) {
        atomic_inc(&c.counters[counter]);
    }
}
SB2_HD bool lane_internal(const Ctx& c, int l) { return (c.lane_flags[l] & LF_INTERNAL) != 0; }
SB2_HD bool state_priority(uint8_t s) { return s >= 'A' && s <= 'Z'; }               // MSLink.h:473
SB2_HD bool state_red(uint8_t s) { return s == 'r' || s == 'u'; }                     // MSLink.h:484
SB2_HD bool state_yellow(uint8_t s) { return s == 'y' || s == 'Y'; }                  // MSLink.h:488
SB2_HD bool state_green(uint8_t s) { return s == 'G' || s == 'g'; }                   // MSLink.h:492

}  // namespace sb2
