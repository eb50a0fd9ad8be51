/* sumob200.h -- C ABI of libsumob200.so, the B200-native replacement of SUMO's per-step vehicle update.
 *
 * The boundary sits where the reference's MSNet::simulationStep calls MSEdgeControl
 * (src/microsim/MSNet.cpp:854-919): one sumob200_step() does what these reference calls do for one simulation step
 *   myBeginOfTimestepEvents->execute (static TLS SwitchCommand)      MSNet.cpp:854, traffic_lights/MSTrafficLightLogic.cpp:68
 *   MSEdgeControl::planMovements / setJunctionApproaches              MSNet.cpp:875,878  (src/microsim/MSEdgeControl.cpp:142,188)
 *   MSEdgeControl::executeMovements                                   MSNet.cpp:881      (MSEdgeControl.cpp:196)
 *   MSEdgeControl::detectCollisions                                   MSNet.cpp:883      (MSEdgeControl.cpp:339)
 *   MSInsertionControl::emitVehicles                                  MSNet.cpp:915      (src/microsim/MSInsertionControl.cpp:127)
 * The scenario crosses the boundary as plain named arrays (the flattened network / vTypes / routes / vehicles described
 * in sumob200_flatfile.h), produced on the reference side by integration/sumo_flatten.cpp from the objects the unchanged
 * NLBuilder built.  No SUMO type, no C++ type and no torch type appears in any signature.
 *
 * All functions return 0 on success and a negative SUMOB200_E* code on failure; sumob200_last_error() gives the text.
 * One engine must be driven by one host thread at a time.  There is no CPU fallback: without a CUDA device every
 * call that computes fails with SUMOB200_ENODEVICE.
 */
#ifndef SUMOB200_H
#define SUMOB200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SUMOB200_EINVAL (-1)      /* bad argument / malformed arrays */
#define SUMOB200_EUNSUPPORTED (-2) /* scenario uses a feature the engine refuses (sublanes, ballistic, multi-lane edges, ...) */
#define SUMOB200_ENODEVICE (-3)   /* no usable CUDA device */
#define SUMOB200_ECUDA (-4)       /* CUDA runtime error */

/* bits of sumob200_step_result.flags: events the engine cannot follow the reference through (emergency stops at a lane end,
 * link states it refuses) or capacity overflows of its fixed-size per-vehicle tables.  A non-zero value means the run diverges
 * from the reference from that step on; the caller must treat it as an error, it is never silently ignored.  Teleports and
 * collisions are EXECUTED on the device like the reference does (MSLane.cpp:2402-2456, 2078-2211, MSVehicleTransfer.cpp) and only
 * counted (teleports, collisions); bits 8 and 16 are kept for ABI stability and never raised. */
#define SUMOB200_F_DRIVE_ITEM_OVERFLOW 1
#define SUMOB200_F_FURTHER_OVERFLOW 2
#define SUMOB200_F_PARTIAL_OVERFLOW 4
#define SUMOB200_F_COLLISION 8
#define SUMOB200_F_TELEPORT 16
#define SUMOB200_F_EMERGENCY_STOP 32
#define SUMOB200_F_NO_CONNECTION 64
#define SUMOB200_F_UNSUPPORTED_LINKSTATE 128
#define SUMOB200_F_LANE_ORDER 256
#define SUMOB200_F_SEARCH_OVERFLOW 512
#define SUMOB200_F_DEFERRED_UNRESOLVED 1024  /* executeMove pass 2 did not converge (sb2_exec.h) */
#define SUMOB200_F_PARTITION_REACH 2048      /* a vehicle's look-ahead left the columns its rank simulates (edges too short for the ghost width) */
#define SUMOB200_F_HALO_OVERFLOW 4096        /* a halo message needed more vehicle records than its capacity */
#define SUMOB200_F_DETECTOR_OVERFLOW 8192    /* more E1 passages in one output interval than the event list holds, or > 4 E1 (or > 8 E2) detectors under one vehicle, or more E2 notifications in a step than 2 per vehicle + 65536 */

typedef struct sumob200_engine sumob200_engine;

/* one named array of the flattened scenario; dtype: 0=u8 1=i32 2=i64 3=f64 4=u64 (sumob200_flatfile.h) */
typedef struct sumob200_array {
    const char* name;
    uint32_t dtype;
    uint32_t elem_size;
    uint64_t count;
    const void* data;     /* host memory, borrowed for the duration of the call */
} sumob200_array;

typedef struct sumob200_step_result {
    int64_t time_ms;          /* simulation time after the last executed step (MSNet::getCurrentTimeStep) */
    int32_t running;          /* vehicles on the road + teleporting ones (MSVehicleControl::getRunningVehicleNo) */
    int32_t departed_total;   /* vehicles inserted so far              (getDepartedVehicleNo) */
    int32_t arrived_total;    /* vehicles that reached their arrival   (getArrivedVehicleNo) */
    int32_t collisions;       /* collisions detected so far */
    int32_t emergency_stops;  /* emergency stops at lane ends so far */
    int32_t teleports;        /* teleports so far: waited longer than time-to-teleport, or collided (MSVehicleControl::getTeleportCount) */
    int32_t flags;            /* OR of SUMOB200_F_* raised so far */
    int32_t halting;          /* on-road vehicles slower than 0.1 m/s     (MSVehicleControl::getHaltingVehicleNo, summary output) */
    int64_t vehicle_updates;  /* sum over the executed steps of vehicles on the road (numerator of SUMO's UPS, MSNet.cpp:959) */
    int32_t teleporting;      /* vehicles currently off the road in MSVehicleTransfer (part of `running`) */
    int32_t discarded_total;  /* vehicles removed before their departure: waited longer than --max-depart-delay (MSInsertionControl::tryInsert) */
} sumob200_step_result;

/* state mirror of one vehicle slot (slot = index in the veh_* arrays = load order); what libsumo::Vehicle getters read
 * from MSVehicle (src/libsumo/Vehicle.cpp getLaneID/getLanePosition/getSpeed) and what MSFCDExport writes */
typedef struct sumob200_vehicle_state {
    int32_t lane;             /* lane index, -1 not yet departed, -2 arrived */
    int32_t route_pos;        /* index of the current route edge */
    double pos;               /* MSVehicle::getPositionOnLane */
    double speed;             /* MSVehicle::getSpeed */
} sumob200_vehicle_state;

/* build an engine on CUDA device `device` from the flattened scenario arrays */
int sumob200_create(const sumob200_array* arrays, int n_arrays, int device, sumob200_engine** out);
/* same, reading a .flat container (sumob200_flatfile.h) */
int sumob200_create_from_file(const char* flat_path, int device, sumob200_engine** out);
void sumob200_destroy(sumob200_engine* e);

/* run n_steps simulation steps; *out (may be NULL) receives the counters after the last one */
int sumob200_step(sumob200_engine* e, int n_steps, sumob200_step_result* out);

/* like sumob200_step, and *elapsed_ms receives the device time of the n_steps steps measured with CUDA events recorded
 * on the engine's own stream (the stream the kernels are launched on) */
int sumob200_step_timed(sumob200_engine* e, int n_steps, sumob200_step_result* out, double* elapsed_ms);

/* copy the state of all vehicle slots to host memory (n = number of loaded vehicles, see sumob200_num_vehicles) */
int sumob200_mirror_all(sumob200_engine* e, sumob200_vehicle_state* out, int capacity);
/* FCD rows of the vehicles currently on the road, ordered by slot: writes up to `capacity` rows, *n = rows available */
typedef struct sumob200_fcd_row {
    int32_t slot;             /* vehicle slot (index into the veh_* arrays; veh_names gives the id) */
    int32_t lane;
    double pos, speed;
    double x, y, angle;       /* MSVehicle::getPosition / computeAngle as navi degrees (MSFCDExport.cpp:119-180) */
} sumob200_fcd_row;
int sumob200_fcd_snapshot(sumob200_engine* e, sumob200_fcd_row* rows, int capacity, int* n);
/* streaming form (north star: "FCD output is streamed via pinned async D2H copies"): _begin compacts the rows on the
 * device and starts their D2H copy into a pinned buffer of the engine on a separate copy stream; the caller may issue the
 * next sumob200_step right away; _end waits for the copy and hands out the pinned rows (valid until the next _begin
 * after the following one: two buffers alternate).  Replaces the per-vehicle loop of MSFCDExport::write
 * (src/microsim/output/MSFCDExport.cpp:98-300), whose XML formatting stays on the host. */
/* host-side formatter: the <timestep> element MSFCDExport::write emits for these rows, byte for byte
 * (ids in std::string order, 2 decimals), formatted on all host threads.  Writes at most `capacity` bytes to buf and the
 * full length to *length (call again with a larger buffer if *length > capacity).  Needs a scenario flattened with ids. */
int sumob200_fcd_format_xml(sumob200_engine* e, const sumob200_fcd_row* rows, int n, int64_t time_ms, char* buf, uint64_t capacity, uint64_t* length);
int sumob200_fcd_begin(sumob200_engine* e);
int sumob200_fcd_end(sumob200_engine* e, const sumob200_fcd_row** rows, int* n);

/* per-trip record of one vehicle slot: the fields MSDevice_Tripinfo::generateOutput writes
 * (src/microsim/devices/MSDevice_Tripinfo.cpp:313-410); times in ms like SUMOTime */
typedef struct sumob200_tripinfo {
    int64_t depart_ms;        /* MSBaseVehicle::getDeparture; -1 = not departed yet */
    int64_t depart_delay_ms;  /* getDepartDelay = departure - depart attribute */
    int64_t arrival_ms;       /* myArrivalTime; -1 = not arrived */
    int64_t duration_ms;      /* arrival (or current time) - departure */
    int64_t waiting_ms;       /* myWaitingTime */
    int64_t time_loss_ms;     /* MSVehicle::getTimeLoss = TIME2STEPS(myTimeLoss) */
    int32_t depart_lane, arrival_lane, waiting_count;
    int32_t teleport_arrived; /* != 0: removed by MSVehicleTransfer as "teleports beyond arrival edge" (tripinfo: vaporized="teleport"); 1 = in
                               * MSVehicleTransfer::add, 2 = in checkInsertions (MSVehicleControl::removePending sees it one step later) */
    double depart_pos, depart_speed, arrival_pos, arrival_speed, route_length, speed_factor;
} sumob200_tripinfo;
/* trip records of all vehicle slots (n = sumob200_num_vehicles) */
int sumob200_tripinfo_all(sumob200_engine* e, sumob200_tripinfo* out, int capacity);

/* the attributes of one <step .../> line of --summary-output (MSNet::writeSummaryOutput, src/microsim/MSNet.cpp:688-730) for
 * the step executed last; `loaded`, `discarded` and `duration` stay with the loader / the host (not engine state) */
typedef struct sumob200_summary {
    int64_t time_ms;          /* time of the step executed last (= the line's time attribute) */
    int32_t inserted;         /* getDepartedVehicleNo */
    int32_t running;          /* getRunningVehicleNo */
    int32_t waiting;          /* MSInsertionControl::getWaitingVehicleNo: depart time reached, no room yet */
    int32_t ended;            /* getEndedVehicleNo */
    int32_t arrived;          /* getArrivedVehicleNo */
    int32_t collisions, teleports;
    int32_t halting;          /* getHaltingVehicleNo */
    int32_t stopped;          /* always 0: <stop>s are outside the path (refused at load) */
    int32_t reserved;
    double mean_waiting_time; /* getTotalDepartureDelay / inserted, -1 if none inserted yet */
    double mean_travel_time;  /* getTotalTravelTime / ended, -1 if none ended yet */
    double mean_speed;        /* getVehicleMeanSpeeds().first, -1 if no vehicle on the road */
    double mean_speed_relative;
} sumob200_summary;
int sumob200_summary_now(sumob200_engine* e, sumob200_summary* out);

/* sumob200_summary_now without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs,
 * same output, same result as sumob200_summary_now, but the kernel's scratch (the per-block partial sums and the reduced result)
 * lies in a file-scope `static __device__` array of the library instead of a lazily allocated buffer: 917,560 bytes that every GPU
 * which loads the library holds until exit.  The call makes no cudaMalloc / cudaFree / allocating library call.
 * Limit: engines with more than SUMOB200_SUMMARY_STATIC_MAX_VEHICLES vehicle slots are rejected with SUMOB200_EINVAL and a message
 * in sumob200_last_error().  All calls on one GPU share the array (it keeps the last call's contents; nothing in it needs to
 * start at zero, the result slot is cleared anyway): calls that overlap on the GPU -- two engines on one device called from two
 * threads -- race and must be serialised by the caller. */
#define SUMOB200_SUMMARY_STATIC_MAX_VEHICLES 2097152
int sumob200_summary_now_static(sumob200_engine* e, sumob200_summary* out);

/* What libsumo::Lane::getLastStepVehicleNumber / getLastStepHaltingNumber / getLastStepMeanSpeed / getLastStepOccupancy /
 * getWaitingTime return (src/libsumo/Lane.cpp:209-272), for every lane at once (the observation vector of RL tooling): out[lane].
 * mean_speed of an empty lane is its speed limit (MSLane::getMeanSpeed); occupancy is the netto occupancy in [0, 1] incl. the part of
 * vehicles lapping in from the next lanes. */
typedef struct sumob200_lane_stats {
    int32_t vehicle_number, halting_number;
    double mean_speed, occupancy, waiting_time;
} sumob200_lane_stats;
int sumob200_lane_stats_all(sumob200_engine* e, sumob200_lane_stats* out, int capacity);

/* E1 induction loops (src/microsim/output/MSInductLoop.cpp; row f4).  Detectors are part of the scenario: optional arrays
 * e1_lane (int32, lane index), e1_pos (float64, metres from the lane's start; negative = from its end, as
 * NLDetectorBuilder::getPositionChecking does), e1_length (float64, optional; 0) -- what <inductionLoop lane= pos= length=/> of an
 * additional file declares.  Vehicles notify the detectors during sumob200_step exactly as MSInductLoop::notifyEnter / notifyMove /
 * notifyLeave do (entry and leave times interpolated inside the step by MSCFModel::passingTime).
 * sumob200_e1_interval = MSInductLoop::writeXMLOutput (MSInductLoop.cpp:456-493) for every detector at once: the aggregates of the
 * interval [begin_ms, end_ms) that ends NOW (end_ms must be the engine's time: call it after the step that completes the period,
 * where MSDetectorControl::writeOutput(step + DELTA_T) fires), written to out[detector]; then MSInductLoop::reset.  The XML line
 * stays with the reference's writer.  Under sumob200_set_partition every rank reports the detectors on its own lanes (the others' records
 * stay empty): merge by owner. */
typedef struct sumob200_e1_interval_t {
    int32_t n_veh_contrib;        /* nVehContrib: vehicles that passed completely */
    int32_t n_veh_entered;        /* nVehEntered */
    double flow;                  /* veh/h */
    double occupancy;             /* % */
    double speed;                 /* mean of the passages' speeds, -1 if none */
    double harmonic_mean_speed;   /* -1 if none */
    double length;                /* mean vehicle length, -1 if none */
} sumob200_e1_interval_t;
int sumob200_e1_count(const sumob200_engine* e);
int sumob200_e1_interval(sumob200_engine* e, int64_t begin_ms, int64_t end_ms, sumob200_e1_interval_t* out, int capacity);
/* The raw passages instead of the aggregates, for a caller that keeps the reference's MSInductLoop objects (the hook of
 * integration/sumo_b200_hook.cpp fills their myVehicleDataCont / myVehiclesOnDet with them, so that the reference's own
 * writeXMLOutput and the libsumo getters work): everything since the last drain / interval call, sorted by (detector, leave time,
 * vehicle), the vehicles that are on a detector right now last per detector with leave_s = SUMOB200_E1_STILL_ON; entered[detector]
 * = vehicles that entered since the last call (may be NULL).  Resets the device-side lists like sumob200_e1_interval. */
#define SUMOB200_E1_STILL_ON (-1e300)
typedef struct sumob200_e1_event {
    int32_t detector;
    int32_t vehicle;          /* slot (load order) */
    double entry_s, leave_s;  /* simulation seconds, interpolated inside the step */
    double vehicle_length;
    int32_t left_early;       /* 1: arrived / teleported while on the detector (MSInductLoop::notifyLeave) */
    int32_t reserved;
} sumob200_e1_event;
int sumob200_e1_drain(sumob200_engine* e, sumob200_e1_event* events, int capacity, int* n, int32_t* entered);

/* E2 lane area detectors (src/microsim/output/MSE2Collector.cpp; row f4).  Part of the scenario: arrays e2_lane (int32), e2_start,
 * e2_length (float64, metres: myStartPos, myDetectorLength of the built detector), e2_jam_time (int64 ms), e2_jam_speed, e2_jam_dist
 * (float64: the timeThreshold / speedThreshold / jamThreshold attributes of <laneAreaDetector>).  One lane per detector.  Vehicles
 * notify the detectors during sumob200_step as MSE2Collector::notifyEnter / notifyMove / notifyLeave do; after every step the engine
 * runs MSE2Collector::detectorUpdate on the device (notifications sorted by distance to the detector's end, sums added in that order,
 * halting durations, jams).  sumob200_e2_interval = MSE2Collector::writeXMLOutput for every detector (the attributes of one
 * <interval .../>, durations in seconds), then MSE2Collector::reset.  Call it after the step that completes the period.
 * Under sumob200_set_partition every rank reports the detectors on its own lanes: merge by owner. */
typedef struct sumob200_e2_interval_t {
    double sampled_seconds;
    int32_t n_veh_entered, n_veh_left, n_veh_seen, max_jam_length_in_vehicles;
    double mean_speed, mean_time_loss, mean_occupancy, max_occupancy, mean_max_jam_length_in_vehicles, mean_max_jam_length_in_meters,
           max_jam_length_in_meters;
    int32_t jam_length_in_vehicles_sum, started_halts;
    double jam_length_in_meters_sum;
    double mean_halting_duration, max_halting_duration, halting_duration_sum;
    double mean_interval_halting_duration, max_interval_halting_duration, interval_halting_duration_sum;
    double mean_vehicle_number;
    int32_t max_vehicle_number, reserved;
} sumob200_e2_interval_t;
int sumob200_e2_count(const sumob200_engine* e);
int sumob200_e2_interval(sumob200_engine* e, sumob200_e2_interval_t* out, int capacity);
/* The raw notifications of the LAST step instead of the aggregates, for a caller that keeps the reference's MSE2Collector objects
 * (the hook of integration/sumo_b200_hook.cpp puts them into myMoveNotifications, so that the reference's own detectorUpdate,
 * writeXMLOutput and the libsumo::LaneArea getters work): one record per (vehicle, detector) that MSE2Collector::notifyMove would
 * have stored in that step, in no particular order; counts[3 * detector + {0, 1, 2}] = vehicles entered / left / seen since the
 * engine was created or sumob200_e2_interval was last called (running totals; may be NULL).  Changes nothing on the device. */
typedef struct sumob200_e2_note {
    int32_t detector;
    int32_t vehicle;                 /* slot (load order) */
    double dist_to_detector_end;
    double time_on_detector, length_on_detector, time_loss, speed;
} sumob200_e2_note;
int sumob200_e2_notes(sumob200_engine* e, sumob200_e2_note* notes, int capacity, int* n, int32_t* counts);

/* edgeData / laneData (src/microsim/output/MSMeanData.cpp:87-253, MSMeanData_Net.cpp:131-246; row f4).  Enabled by the scenario
 * scalar md_enabled = 1 (optional md_halt_speed = haltingSpeedThreshold, default 0.1): every vehicle then reports the part of each
 * step it spends on each non-internal lane exactly as MSMeanData::MeanDataValues::notifyMove does, and the engine accumulates per
 * lane the raw members of MSMeanData_Net::MSLaneMeanDataValues.  sumob200_meandata_interval copies them out for ALL lanes
 * (out[lane]; internal lanes stay zero) and starts a new interval (MSLaneMeanDataValues::reset).  The derived attributes of the
 * <edge .../> line (density, traveltime, speed, flow, ...) are MSLaneMeanDataValues::write's business and stay with the reference's
 * writer; edges with several lanes are summed by MSLaneMeanDataValues::addTo there.  The sums are accumulated with atomics: their
 * last bits depend on the order of additions (the reference's differ too: it adds in lane execution order). */
typedef struct sumob200_lane_meandata {
    double sample_seconds, travelled_distance, wait_seconds, time_loss;
    double front_sample_seconds, front_travelled_distance, veh_length_sum, occupation_sum;
    double minimal_vehicle_length;   /* < 0: no vehicle seen (INVALID_DOUBLE) */
    int32_t departed, arrived, entered, left, teleported, reserved;
} sumob200_lane_meandata;
int sumob200_meandata_interval(sumob200_engine* e, sumob200_lane_meandata* out, int capacity);

/* accumulated emissions of every vehicle slot since its departure (MSDevice_Emissions::myEmissions, what tripinfo's
 * <emissions .../> and --emission-output report; src/microsim/devices/MSDevice_Emissions.cpp:122-135): out[p * n + slot],
 * p = 0 CO2, 1 CO, 2 HC, 3 fuel, 4 NOx, 5 PMx, 6 electricity [mg resp. ml resp. Wh], n = sumob200_num_vehicles.
 * capacity = number of doubles in out (>= 7 * n).  SUMOB200_EUNSUPPORTED if no vehicle carries the emission device. */
int sumob200_emissions(sumob200_engine* e, double* out, int64_t capacity);

/* TraCI-style traffic light control between two steps.  tls = index of the program in the flattened scenario (tl_names).
 * _set_phase = libsumo::TrafficLight::setPhase (src/libsumo/TrafficLight.cpp:823-832): switch to `phase` now, its full duration
 * starts now (MSSimpleTrafficLightLogic::changeStepAndDuration, MSSimpleTrafficLightLogic.cpp:294-309); a phase index out of
 * range is SUMOB200_EINVAL with TraCI's message.  _get_phase = getPhase / getNextSwitch. */
int sumob200_num_tls(const sumob200_engine* e);
int sumob200_tls_set_phase(sumob200_engine* e, int tls, int phase);
int sumob200_tls_get_phase(sumob200_engine* e, int tls, int* phase, int64_t* next_switch_ms);

/* libsumo::Vehicle::setSpeed (src/libsumo/Vehicle.cpp:1924-1938) for n vehicle slots between two steps: from the next step on
 * the vehicle drives speeds[i] as far as its safe speed and its acceleration / deceleration bounds allow (default speed mode 31:
 * MSVehicle::Influencer::influenceSpeed, MSVehicle.cpp:493-520, applied in planMove :2313-2326 and executeMove :4025-4056);
 * speeds[i] < 0 hands the vehicle back to the car-following model.  An unknown slot is SUMOB200_EINVAL. */
int sumob200_vehicles_set_speed(sumob200_engine* e, const int32_t* slots, const double* speeds, int n);

/* ---- multi-GPU: one connected network cut into strips of junction columns, one engine (process, GPU) per strip ----------------
 * Mirrors the per-lane tasking of MSEdgeControl::planMovements / executeMovements (src/microsim/MSEdgeControl.cpp:150-167, 205-219)
 * across GPUs; design in sumo_b200/csrc/sb2_halo.h.  Every engine is created from the WHOLE scenario (same arrays, same slot
 * numbers) and then told its part:
 *   lane_vis[n_lanes]  3 = own lane (simulated, authoritative), 2 = ghost (first column of a neighbour: simulated redundantly),
 *                      1 = halo (second column: read only), 0 = not visible
 *   exp_lanes[side] / imp_lanes[side]   side 0 = neighbour towards lower columns, 1 = higher: the lanes whose final state this rank
 *                      sends (its own two border columns) / receives (the neighbour's); the neighbour's export list is this rank's
 *                      import list, element for element;  n = 0 for a side without neighbour
 *   capacity           vehicle records per message (the same on all ranks); an overflow raises SUMOB200_F_HALO_OVERFLOW
 * After every sumob200_step(e, 1, ..) the caller runs, per side with a neighbour: sumob200_halo_pack into a device buffer of
 * sumob200_halo_message_bytes(e, side, 1) bytes, transports it (ncclSend/ncclRecv, cudaMemcpyPeer ... on sumob200_stream's order),
 * and sumob200_halo_unpack on the buffer received from that neighbour.  The three fleet aggregates behind MSLane::getMaximumBrakeDist
 * (max speedFactor, -min decel, max minGap over all departed vehicles, MSVehicleControl.cpp:199-216) are network-wide:
 * sumob200_fleet_aggregates returns their device address for a MAX all-reduce per step (the second one is stored negated
 * only inside that call sequence: negate, all-reduce MAX, negate). */
int sumob200_set_partition(sumob200_engine* e, const uint8_t* lane_vis, const int32_t* const* exp_lanes, const int32_t* n_exp,
                           const int32_t* const* imp_lanes, const int32_t* n_imp, int capacity);
int64_t sumob200_halo_message_bytes(const sumob200_engine* e, int side, int outgoing);
/* max_records > 0 limits the vehicle records of this message (<= capacity): lets both ends transport a prefix of the buffer whose
 * size they derive from the previous step's count; more records than that raise SUMOB200_F_HALO_OVERFLOW */
int sumob200_halo_pack(sumob200_engine* e, int side, void* device_message, int max_records);
int sumob200_halo_unpack(sumob200_engine* e, int side, const void* device_message);
void* sumob200_stream(sumob200_engine* e);            /* the cudaStream_t every kernel of the engine is launched on */
double* sumob200_fleet_aggregates(sumob200_engine* e);  /* device pointer, 3 doubles */

int sumob200_num_vehicles(const sumob200_engine* e);
int sumob200_num_lanes(const sumob200_engine* e);
int64_t sumob200_time_ms(const sumob200_engine* e);
/* supply the [0,1) dawdle draws of the next step (one per vehicle slot), e.g. draws recorded from the reference; n must
 * equal sumob200_num_vehicles.  After the call every step consumes draws[] until sumob200_set_random_mode resets it. */
int sumob200_set_random_draws(sumob200_engine* e, const double* draws, int n);
/* 0 = no draws (only valid if all sigma == 0), 1 = counter-based generator seeded by --seed */
int sumob200_set_random_mode(sumob200_engine* e, int mode);
/* number of kernel launches issued by the engine so far (bench.py reports it as gpu_launches) */
int64_t sumob200_kernel_launches(const sumob200_engine* e);
/* time spent in the kernels of the step pipeline, measured with CUDA events on the engine's stream: per phase sums in ms
 * since the last call with reset != 0.  phases: 0 pre (tls, lane summary) 1 plan 2 execute 3 hand-off 4 insert */
int sumob200_phase_times(sumob200_engine* e, double* ms, int n_phases, int reset);
int sumob200_set_profiling(sumob200_engine* e, int enabled);
/* memory layout tuning: every `steps` steps (0 = never, the default) the engine renumbers the on-road vehicles in lane order, so
 * that vehicles that are neighbours on the road are neighbours in HBM.  Results are unchanged; the vehicle indices of this API
 * ("slot") stay the load-order indices.  With it on, sumob200_fcd_* rows come in memory order instead of slot order.  Ignored in
 * partitioned runs (all ranks share one numbering). */
int sumob200_set_reslot_period(sumob200_engine* e, int steps);

const char* sumob200_last_error(void);
const char* sumob200_version(void);

#ifdef __cplusplus
}
#endif
#endif
