/* sumo_b200_hook.cpp -- REFERENCE-SIDE binding of the engine: what makes `sumo` drive libsumob200.so instead of MSEdgeControl.
 *
 * Built by oracle/build_ref.sh (target sumo_b200) into oracle/_ref/sumo_b200 = the reference's own objects, with two changes:
 *   1. MSNet.cpp is compiled from an OVERLAY COPY generated at build time (oracle/_ref/gen/MSNet_b200.cpp, never committed):
 *      the reference's file with one line inserted at the top of MSNet::simulationStep (src/microsim/MSNet.cpp:799)
 *          if (sumob200_hook_step(this)) { return; }
 *      -- the patch of INTEGRATION.md, applied by `sed`.
 *   2. this file replaces sumo_main.cpp: the same start-up sequence (XMLSubSys, OptionsIO, NLBuilder::init, MSNet::simulate,
 *      closeSimulation) plus the option `--b200.device N`.
 * Everything else -- NLBuilder network / route loading, the option parser, MSFCDExport and the other output writers,
 * libsumo -- is the reference's unchanged code.
 *
 * With `--b200.device N` the first step flattens the loaded scenario IN MEMORY (the walk of integration/sumo_flatten.cpp over
 * MSEdge / MSLane / MSLink / MSTrafficLightLogic / MSVehicleType / SUMOVehicle) and hands the arrays to sumob200_create; every
 * step then is: sumob200_step(1) on the GPU, the vehicle states mirrored back into the reference's MSVehicle objects (lane,
 * position, speed, angle, on-road flag) and MSVehicleControl's counters, MSNet::writeOutput (so --fcd-output is written by the
 * reference's own MSFCDExport from those objects), time += DELTA_T.  libsumo getters (Vehicle::getSpeed, getLaneID,
 * getLanePosition ...) read the same mirrored objects.
 *
 * This is integration / test code on the reference's side of the boundary: it reaches MSVehicle's state through
 * `#define private public` (sumo_flatten.cpp), which a maintainer would replace by a friend declaration or setters.
 *
 * `--libsumo-check K` instead of simulate(): K steps through libsumo::Simulation::step() and after each one the ids, lanes,
 * positions and speeds as libsumo::Vehicle reports them, printed to stdout (tests/test_boundary.py compares the output of the
 * stock run with the engine run).
 */
#include <config.h>
#include <algorithm>
#include <bitset>
#include <list>
#include <map>
#include <random>
#include <set>
#include <sstream>
#include <string>
#include <vector>
#include <memory>
#include <iostream>
#include <fstream>
#include <functional>
#include <cassert>
#include <limits>
#include <utils/common/StdDefs.h>
#include <utils/common/SUMOTime.h>
#include <utils/geom/Position.h>
#include <utils/geom/PositionVector.h>
// write access to the few members the mirror sets (MSVehicle: lane, state, angle, on-road flag; MSBaseVehicle: route position,
// departure; MSVehicleControl: counters; MSNet: clock) -- see the header comment
#define private public
#define protected public
#include <microsim/MSBaseVehicle.h>
#include <microsim/MSVehicle.h>
#include <microsim/MSVehicleControl.h>
#include <microsim/MSNet.h>
#include <microsim/MSInsertionControl.h>
#include <microsim/devices/MSDevice_Tripinfo.h>
#include <microsim/output/MSMeanData.h>
#include <microsim/output/MSMeanData_Net.h>
#include <microsim/output/MSDetectorControl.h>
#include <microsim/output/MSInductLoop.h>
#include <microsim/output/MSE2Collector.h>
#undef private
#undef protected
#define SUMOB200_NO_MAIN
#include "sumo_flatten.cpp"

#include <libsumo/Simulation.h>
#include <utils/geom/GeomHelper.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "../include/sumob200.h"

namespace {

int gDevice = -1;

struct Hook {
    sumob200_engine* eng = nullptr;
    std::vector<MSVehicle*> vehOfSlot;
    std::vector<MSLane*> laneOfIndex;
    std::vector<sumob200_vehicle_state> state;
    std::vector<sumob200_fcd_row> rows;
    std::vector<char> wasOnRoad;                     // 0 not departed, 1 departed, 2 removed (arrived)
    std::vector<sumob200_tripinfo> trips;
    std::vector<SUMOTime> departBucket;               // MSVehicleContainer's bucket of every slot (depart rounded up to a step)
    std::vector<MSInductLoop*> loops;                 // the E1 detectors in the order the flatten walk exported them (id order)
    std::vector<sumob200_e1_event> e1Events;
    std::vector<MSE2Collector*> areas;                // the E2 detectors, same order as the engine's
    std::vector<sumob200_e2_note> e2Notes;
    std::vector<int32_t> e2Counts, e2CountsLast;
    std::vector<int32_t> e1Entered;
    std::vector<std::vector<MSMeanData_Net::MSLaneMeanDataValues*> > mdOfLane;   // per lane index: its values object in every <edgeData> / <laneData>
    std::vector<sumob200_lane_meandata> mdBuf;
    int lastDeparted = 0, lastArrived = 0;

    static Hook* create(MSNet* net) {
        sb2::FlatFile ff;
        flattenTo(net, ff);
        std::vector<sumob200_array> arrays;
        for (auto& kv : ff.arrays) {
            sumob200_array a;
            a.name = kv.first.c_str();
            a.dtype = kv.second.dtype;
            a.elem_size = kv.second.elemSize;
            a.count = kv.second.count;
            a.data = kv.second.bytes.data();
            arrays.push_back(a);
        }
        Hook* h = new Hook();
        if (sumob200_create(arrays.data(), (int)arrays.size(), gDevice, &h->eng) != 0) {
            throw ProcessError(std::string("sumob200_create failed: ") + sumob200_last_error());
        }
        for (const MSEdge* e : MSEdge::getAllEdges()) {
            for (MSLane* l : e->getLanes()) {
                h->laneOfIndex.push_back(l);
            }
        }
        std::map<int, MSVehicle*> byNumId;
        MSVehicleControl& vc = net->getVehicleControl();
        for (auto it = vc.loadedVehBegin(); it != vc.loadedVehEnd(); ++it) {
            MSVehicle* v = dynamic_cast<MSVehicle*>(it->second);
            if (v != nullptr) {
                byNumId[v->getNumericalID()] = v;
            }
        }
        const int32_t* numid = ff.ptr<int32_t>("veh_numid");
        const int nV = (int)ff.count("veh_numid");
        for (int s = 0; s < nV; s++) {
            h->vehOfSlot.push_back(byNumId[numid[s]]);
        }
        for (const auto& item : net->getDetectorControl().getTypedDetectors(SUMO_TAG_INDUCTION_LOOP)) {
            h->loops.push_back(dynamic_cast<MSInductLoop*>(item.second));
        }
        if ((int)h->loops.size() != sumob200_e1_count(h->eng)) {
            throw ProcessError("sumob200: detector lists of the reference and the engine differ");
        }
        for (const auto& item : net->getDetectorControl().getTypedDetectors(SUMO_TAG_LANE_AREA_DETECTOR)) {
            h->areas.push_back(dynamic_cast<MSE2Collector*>(item.second));
        }
        if ((int)h->areas.size() != sumob200_e2_count(h->eng)) {
            throw ProcessError("sumob200: lane area detector lists of the reference and the engine differ");
        }
        h->e2Notes.resize(h->areas.empty() ? 1 : 2 * (size_t)nV + 65536);
        h->e2Counts.assign(3 * h->areas.size() + 1, 0);
        h->e2CountsLast.assign(3 * h->areas.size() + 1, 0);
        {
            std::map<const MSLane*, int> indexOfLane;
            for (size_t i = 0; i < h->laneOfIndex.size(); i++) {
                indexOfLane[h->laneOfIndex[i]] = (int)i;
            }
            bool any = false;
            h->mdOfLane.resize(h->laneOfIndex.size());
            for (const auto& item : net->getDetectorControl().getMeanData()) {
                for (MSMeanData* md : item.second) {
                    for (auto& edgeValues : md->myMeasures) {
                        for (MSMeanData::MeanDataValues* v : edgeValues) {
                            MSMeanData_Net::MSLaneMeanDataValues* lv = dynamic_cast<MSMeanData_Net::MSLaneMeanDataValues*>(v);
                            if (lv == nullptr || lv->getLane() == nullptr) {
                                throw ProcessError("sumob200: unexpected mean data container in '" + item.first + "'");
                            }
                            h->mdOfLane[indexOfLane[lv->getLane()]].push_back(lv);
                            any = true;
                        }
                    }
                }
            }
            if (any) {
                h->mdBuf.resize(h->laneOfIndex.size());
            } else {
                h->mdOfLane.clear();
            }
        }
        h->e1Entered.resize(h->loops.size() + 1);
        h->e1Events.resize(4 * (size_t)nV + 65536);
        h->state.resize(nV > 0 ? nV : 1);
        h->rows.resize(nV > 0 ? nV : 1);
        h->wasOnRoad.assign(nV, 0);
        h->trips.resize(nV > 0 ? nV : 1);
        for (int s = 0; s < nV; s++) {
            const SUMOTime d = h->vehOfSlot[s]->getParameter().depart;
            const SUMOTime delay = d % DELTA_T;
            h->departBucket.push_back(d + (delay == 0 ? 0 : DELTA_T - delay));
        }
        WRITE_MESSAGE("sumob200: " + std::string(sumob200_version()) + "; " + toString(nV) + " vehicles, " + toString(h->laneOfIndex.size()) + " lanes on CUDA device " + toString(gDevice));
        return h;
    }

    void step(MSNet* net) {
        for (const MSEdge* e : MSEdge::getAllEdges()) {
            if (e->isVaporizing()) {   // (<vaporizer> starts as an event, nothing at load time tells)
                throw ProcessError("sumob200: vaporizers are not supported (edge '" + e->getID() + "')");
            }
        }
        sumob200_step_result r;
        if (sumob200_step(eng, 1, &r) != 0) {
            throw ProcessError(std::string("sumob200_step failed: ") + sumob200_last_error());
        }
        if (r.flags != 0) {
            throw ProcessError("sumob200: the engine raised flags " + toString(r.flags) + " (the run left the supported path)");
        }
        const int nV = (int)vehOfSlot.size();
        int nRows = 0;
        if (sumob200_mirror_all(eng, state.data(), nV) != 0 || sumob200_fcd_snapshot(eng, rows.data(), nV, &nRows) != 0) {
            throw ProcessError(std::string("sumob200 mirror failed: ") + sumob200_last_error());
        }
        // ---- mirror: what libsumo::Vehicle getters and MSFCDExport read from MSVehicle
        for (int i = 0; i < nRows; i++) {
            const sumob200_fcd_row& row = rows[i];
            MSVehicle* v = vehOfSlot[row.slot];
            if (!wasOnRoad[row.slot]) {
                wasOnRoad[row.slot] = 1;
                v->myDeparture = net->getCurrentTimeStep();   // inserted in this step
                v->myDepartPos = row.pos;                     // (a vehicle does not move in the step of its insertion)
                net->getVehicleControl().vehicleDeparted(*v); // ++running, total departure delay (summary output)
            }
            v->myLane = laneOfIndex[row.lane];
            v->myAmOnNet = true;
            v->myState.myPos = row.pos;
            v->myState.mySpeed = row.speed;
            v->myState.myBackPos = row.pos - v->getVehicleType().getLength();
            v->myCachedPosition = Position::INVALID;            // MSVehicle::getPosition recomputes it from lane + position
            v->myAngle = DEG2RAD(90. - row.angle);              // inverse of GeomHelper::naviDegree
            v->myCurrEdge = v->getRoute().begin() + state[row.slot].route_pos;
        }
        // ---- arrivals: the trip record into the vehicle's MSDevice_Tripinfo, then the reference's own removal
        // (MSVehicleControl::scheduleVehicleRemoval / removePending: tripinfo line, total travel time, counters, delete)
        MSVehicleControl& vcr = net->getVehicleControl();
        bool tripsFetched = false;
        std::vector<int> removed;
        std::vector<SUMOVehicle*> pendingEmits;
        const SUMOTime now = net->getCurrentTimeStep();
        for (int s = 0; s < nV; s++) {
            MSVehicle* v = vehOfSlot[s];
            if (v == nullptr) {
                continue;
            }
            if (wasOnRoad[s] == 0) {
                if (departBucket[s] <= now) {
                    pendingEmits.push_back(v);   // depart time reached, no room yet: MSInsertionControl::myPendingEmits
                }
                continue;
            }
            if (state[s].lane >= 0) {
                continue;
            }
            v->myAmOnNet = false;   // arrived or teleporting: not visible to the output writers
            if (state[s].lane != -2) {
                continue;
            }
            if (!tripsFetched) {
                if (sumob200_tripinfo_all(eng, trips.data(), nV) != 0) {
                    throw ProcessError(std::string("sumob200_tripinfo_all failed: ") + sumob200_last_error());
                }
                tripsFetched = true;
            }
            const sumob200_tripinfo& tr = trips[s];
            if (tr.teleport_arrived == 2 && tr.arrival_ms == now) {
                continue;   // scheduled by MSVehicleTransfer::checkInsertions: removePending runs in the next step's executeMovements
            }
            v->myDeparture = tr.depart_ms;
            v->myDepartPos = tr.depart_pos;
            v->myTimeLoss = (double)tr.time_loss_ms / 1000.;
            v->myCurrEdge = v->getRoute().end() - 1;
            v->myState.mySpeed = tr.arrival_speed;
            MSDevice_Tripinfo* d = static_cast<MSDevice_Tripinfo*>(v->getDevice(typeid(MSDevice_Tripinfo)));
            if (d != nullptr) {
                d->myDepartLane = laneOfIndex[tr.depart_lane]->getID();
                d->myDepartSpeed = tr.depart_speed;
                d->myWaitingTime = tr.waiting_ms;
                d->myWaitingCount = tr.waiting_count;
                d->myStoppingTime = 0;
                d->myArrivalTime = tr.arrival_ms;
                d->myArrivalLane = tr.arrival_lane >= 0 ? laneOfIndex[tr.arrival_lane]->getID() : "";
                d->myArrivalPos = tr.arrival_pos;
                d->myArrivalSpeed = tr.arrival_speed;
                d->myArrivalReason = tr.teleport_arrived ? MSMoveReminder::NOTIFICATION_TELEPORT_ARRIVED : MSMoveReminder::NOTIFICATION_ARRIVED;
                d->myRouteLength = tr.route_length - tr.arrival_pos;
            }
            vcr.scheduleVehicleRemoval(v, false);
            removed.push_back(s);
        }

        // ---- E1 detectors: this step's passages into the reference's own MSInductLoop objects (their writeXMLOutput, called by
        // MSDetectorControl::writeOutput below, and the libsumo::InductionLoop getters then work unchanged)
        if (!loops.empty()) {
            int nEv = 0;
            if (sumob200_e1_drain(eng, e1Events.data(), (int)e1Events.size(), &nEv, e1Entered.data()) != 0) {
                throw ProcessError(std::string("sumob200_e1_drain failed: ") + sumob200_last_error());
            }
            for (size_t d = 0; d < loops.size(); d++) {
                loops[d]->myVehiclesOnDet.clear();
                loops[d]->myEnteredVehicleNumber += e1Entered[d];
            }
            for (int i = 0; i < nEv; i++) {
                const sumob200_e1_event& ev = e1Events[i];
                MSInductLoop* loop = loops[ev.detector];
                MSVehicle* v = vehOfSlot[ev.vehicle];
                if (ev.leave_s == SUMOB200_E1_STILL_ON) {
                    loop->myVehiclesOnDet[v] = ev.entry_s;
                } else if (ev.left_early) {
                    loop->myVehicleDataCont.push_back(MSInductLoop::VehicleData(*v, ev.entry_s, ev.leave_s, true));
                    loop->myLastLeaveTime = ev.leave_s;
                } else {
                    loop->myVehicleDataCont.push_back(MSInductLoop::VehicleData(*v, ev.entry_s, ev.leave_s, false, loop->myEndPosition - loop->myPosition));
                    loop->myLastLeaveTime = ev.leave_s;
                }
            }
        }
        // ---- E2 detectors: this step's notifications into the reference's MSE2Collector::myMoveNotifications; its own detectorUpdate
        // (MSDetectorControl::updateDetectors in writeOutput below) sorts and aggregates them, its writeXMLOutput prints them
        if (!areas.empty()) {
            int nNotes = 0;
            if (sumob200_e2_notes(eng, e2Notes.data(), (int)e2Notes.size(), &nNotes, e2Counts.data()) != 0) {
                throw ProcessError(std::string("sumob200_e2_notes failed: ") + sumob200_last_error());
            }
            for (int i = 0; i < nNotes; i++) {
                const sumob200_e2_note& n = e2Notes[i];
                areas[n.detector]->myMoveNotifications.push_back(new MSE2Collector::MoveNotificationInfo(vehOfSlot[n.vehicle]->getID(), 0., 0., n.speed, 0.,
                        n.dist_to_detector_end, n.time_on_detector, n.length_on_detector, n.time_loss, 0., true));
            }
            for (size_t d = 0; d < areas.size(); d++) {
                areas[d]->myNumberOfEnteredVehicles += e2Counts[3 * d] - e2CountsLast[3 * d];
                areas[d]->myNumberOfLeftVehicles += e2Counts[3 * d + 1] - e2CountsLast[3 * d + 1];
                areas[d]->myNumberOfSeenVehicles += e2Counts[3 * d + 2] - e2CountsLast[3 * d + 2];
            }
            e2CountsLast = e2Counts;
        }
        // ---- edgeData / laneData: this step's contributions added to the reference's MSLaneMeanDataValues objects, whose own
        // write() / reset() (MSMeanData::writeXMLOutput, driven by MSDetectorControl::writeOutput below) then do the rest
        if (!mdOfLane.empty()) {
            if (sumob200_meandata_interval(eng, mdBuf.data(), (int)mdBuf.size()) != 0) {
                throw ProcessError(std::string("sumob200_meandata_interval failed: ") + sumob200_last_error());
            }
            for (size_t l = 0; l < mdOfLane.size(); l++) {
                const sumob200_lane_meandata& m = mdBuf[l];
                if (m.sample_seconds == 0 && m.departed == 0 && m.arrived == 0 && m.entered == 0 && m.left == 0) {
                    continue;
                }
                for (MSMeanData_Net::MSLaneMeanDataValues* lv : mdOfLane[l]) {
                    lv->sampleSeconds += m.sample_seconds;
                    lv->travelledDistance += m.travelled_distance;
                    lv->waitSeconds += m.wait_seconds;
                    lv->timeLoss += m.time_loss;
                    lv->frontSampleSeconds += m.front_sample_seconds;
                    lv->frontTravelledDistance += m.front_travelled_distance;
                    lv->vehLengthSum += m.veh_length_sum;
                    lv->occupationSum += m.occupation_sum;
                    if (m.minimal_vehicle_length >= 0) {
                        lv->minimalVehicleLength = lv->minimalVehicleLength == INVALID_DOUBLE ? m.minimal_vehicle_length : MIN2(lv->minimalVehicleLength, m.minimal_vehicle_length);
                    }
                    lv->nVehDeparted += m.departed;
                    lv->nVehArrived += m.arrived;
                    lv->nVehEntered += m.entered;
                    lv->nVehLeft += m.left;
                    lv->nVehTeleported += m.teleported;
                }
            }
        }
        // ---- the arrived vehicles leave the reference's containers now (after the detectors have copied what they need of them)
        for (const int s : removed) {
            vehOfSlot[s] = nullptr;
            wasOnRoad[s] = 2;
        }
        vcr.removePending();
        net->getInsertionControl().myPendingEmits = pendingEmits;
        // ---- MSVehicleControl counters (summary output, MSNet::simulationState's termination test)
        MSVehicleControl& vc = net->getVehicleControl();
        // (myRunningVehNo / myEndedVehNo / the totals are kept by vehicleDeparted and removePending above)
        vc.myCollisions = r.collisions;
        vc.myTeleportsJam = r.teleports;
        // ---- the rest of MSNet::simulationStep / postMoveStep (MSNet.cpp:919-964): outputs, clock
        net->writeOutput();
        net->myStep += DELTA_T;
    }
};

Hook* gHook = nullptr;

}  // namespace

// called from the inserted line of the overlay MSNet::simulationStep; false = not active, run the reference's own step
bool sumob200_hook_step(MSNet* net) {
    if (gDevice < 0) {
        return false;
    }
    if (gHook == nullptr) {
        gHook = Hook::create(net);
    }
    gHook->step(net);
    return true;
}

int
main(int argc, char** argv) {
    std::vector<char*> args;
    args.push_back(argv[0]);
    int libsumoCheck = -1;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        if (a == "--b200.device" && i + 1 < argc) {
            gDevice = atoi(argv[++i]);
        } else if (a == "--libsumo-check" && i + 1 < argc) {
            libsumoCheck = atoi(argv[++i]);
        } else {
            args.push_back(argv[i]);
        }
    }
    // the engine needs the whole demand at create (incremental route loading = sumob200_add_routes, not built): load it all at start
    static char routeSteps[] = "--route-steps";
    static char zero[] = "0";
    args.push_back(routeSteps);
    args.push_back(zero);
    OptionsCont& oc = OptionsCont::getOptions();
    oc.setApplicationName("sumo", "Eclipse SUMO sumo " VERSION_STRING);
    gSimulation = true;
    int ret = 0;
    MSNet* net = nullptr;
    try {
        XMLSubSys::init();
        OptionsIO::setArgs((int)args.size(), args.data());
        net = NLBuilder::init();
        if (net == nullptr) {
            return 1;
        }
        const SUMOTime begin = string2time(oc.getString("begin"));
        const SUMOTime end = string2time(oc.getString("end"));
        if (libsumoCheck >= 0) {
            // libsumo-style stepping: Simulation::step -> MSNet::simulationStep -> (hook) ; getters read MSVehicle
            for (int k = 0; k < libsumoCheck; k++) {
                libsumo::Simulation::step();
                std::vector<std::string> ids = libsumo::Vehicle::getIDList();
                std::sort(ids.begin(), ids.end());
                printf("t=%.2f n=%d", libsumo::Simulation::getTime(), (int)ids.size());
                for (const std::string& id : ids) {
                    printf(" %s:%s:%.6f:%.6f", id.c_str(), libsumo::Vehicle::getLaneID(id).c_str(), libsumo::Vehicle::getLanePosition(id), libsumo::Vehicle::getSpeed(id));
                }
                printf("\n");
            }
            net->closeSimulation(begin, "libsumo check done");
        } else {
            MSNet::SimulationState state = net->simulate(begin, end);
            (void)state;
        }
        delete net;
        OutputDevice::closeAll();
        MsgHandler::cleanupOnEnd();
    } catch (const ProcessError& e) {
        if (std::string(e.what()) != std::string("Process Error") && std::string(e.what()) != std::string("")) {
            WRITE_ERROR(e.what());
        }
        ret = 1;
    }
    if (gHook != nullptr && gHook->eng != nullptr) {
        sumob200_destroy(gHook->eng);
    }
    SystemFrame::close();
    return ret;
}
