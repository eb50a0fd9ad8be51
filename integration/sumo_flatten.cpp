/* sumo_flatten -- REFERENCE-SIDE integration shim (not part of libsumob200.so).
 *
 * This is the piece a SUMO maintainer adds on their side of the drop-in boundary:
 * it is compiled against the reference headers and linked with the reference's own loader objects (NLBuilder, NLHandler,
 * MSRouteHandler, ...), lets the UNCHANGED loader build MSNet from net/route/sumocfg input, then walks
 * MSEdge/MSLane/MSLink/MSJunctionLogic/MSTrafficLightLogic/MSVehicleType/MSRoute/SUMOVehicle and writes the flattened
 * arrays of include/sumob200_flatfile.h.  The engine never sees MS* objects.
 *
 * It doubles as the parity harness: with --trace-out it then runs the reference's own MSNet::simulationStep and
 * records, per step, every on-road vehicle's (numerical id, lane, pos, speed) as raw fp64 -- the ground truth the
 * "-m gpu" parity tests compare the CUDA engine with bit for bit.
 *
 * usage: sumo_flatten --flat-out X.flat [--trace-out X.trace] <any sumo options: -n net -r routes / -c cfg ...>
 *
 * Build: oracle/build_ref.sh flatten   (objects come from the reference sources under /root/reference, outputs into
 * oracle/_ref/, never committed).
 */
#include <config.h>
#include <version.h>
// The HBEFA coefficient tables (HelpersHBEFA3/4::myFunctionParameter) and the emission device's accumulators
// (MSDevice_Emissions::myEmissions) have no public accessor; a maintainer's in-tree shim would be a friend.  Here, outside
// the tree and without touching it, the three headers are read with private members visible.
#include <string>
#include <vector>
#include <map>
#include <set>
#include <bitset>
#include <sstream>
#define private public
#include <utils/emissions/PollutantsInterface.h>
#include <utils/emissions/HelpersHBEFA3.h>
#include <utils/emissions/HelpersHBEFA4.h>
#include <microsim/devices/MSDevice_Emissions.h>
#include <microsim/MSInsertionControl.h>   // (expandFlows files myPendingEmits back into myAllVeh)
#define protected public
#include <microsim/output/MSDetectorControl.h>   // (myIntervals: which detectors write periodically)
#include <microsim/output/MSE2Collector.h>       // (jam thresholds, lane list, nextEdges filter)
#undef protected
#undef private
#include <utils/emissions/EnergyParams.h>
#include <microsim/MSLink.h>
#include <microsim/MSNet.h>
#include <microsim/MSEdge.h>
#include <microsim/MSLane.h>
#include <microsim/MSJunction.h>
#include <microsim/MSJunctionLogic.h>
#include <microsim/MSRoute.h>
#include <microsim/MSGlobals.h>
#include <microsim/MSVehicle.h>
#include <microsim/MSVehicleType.h>
#include <microsim/MSVehicleControl.h>
#include <microsim/MSInsertionControl.h>
#include <microsim/devices/MSRoutingEngine.h>
#include <microsim/cfmodels/MSCFModel.h>
#include <microsim/traffic_lights/MSTLLogicControl.h>
#include <microsim/traffic_lights/MSTrafficLightLogic.h>
#include <microsim/traffic_lights/MSSimpleTrafficLightLogic.h>
#include <netload/NLBuilder.h>
#include <utils/common/MsgHandler.h>
#include <utils/common/RandHelper.h>
#include <utils/common/SystemFrame.h>
#include <utils/common/StringUtils.h>
#include <utils/options/OptionsCont.h>
#include <utils/options/OptionsIO.h>
#include <utils/xml/XMLSubSys.h>
#include <utils/iodevices/OutputDevice.h>
#include <utils/iodevices/OutputDevice_String.h>
#include <utils/common/StringTokenizer.h>
#include <microsim/MSEdgeControl.h>
#include <microsim/output/MSDetectorControl.h>
#include <microsim/output/MSInductLoop.h>
#include <microsim/output/MSMeanData_Net.h>
#include <microsim/trigger/MSLaneSpeedTrigger.h>
#include <microsim/trigger/MSTriggeredRerouter.h>
#include <microsim/trigger/MSCalibrator.h>
#include <random>
#include <list>
#include <set>
#include <traci-server/TraCIServer.h>
#include <libsumo/TrafficLight.h>
#include <libsumo/Vehicle.h>

#include <algorithm>
#include <map>
#include <string>
#include <vector>

#include "../include/sumob200_flatfile.h"

namespace {

// MSLink::myRadius / myAmCont / myAmContOff / myHavePedestrianCrossingFoe have no getters (MSLink.h:728ff); read them through
// the explicit-instantiation access idiom instead of patching the reference header.
template<typename Tag, typename Tag::type M> struct Rob {
    friend typename Tag::type get(Tag) {
        return M;
    }
};
struct RadiusTag {
    typedef double MSLink::* type;
    friend type get(RadiusTag);
};
template struct Rob<RadiusTag, &MSLink::myRadius>;
struct AmContTag {
    typedef bool MSLink::* type;
    friend type get(AmContTag);
};
template struct Rob<AmContTag, &MSLink::myAmCont>;
struct AmContOffTag {
    typedef bool MSLink::* type;
    friend type get(AmContOffTag);
};
template struct Rob<AmContOffTag, &MSLink::myAmContOff>;
struct PedFoeTag {
    typedef bool MSLink::* type;
    friend type get(PedFoeTag);
};
template struct Rob<PedFoeTag, &MSLink::myHavePedestrianCrossingFoe>;
struct MdHaltTag {
    typedef const double MSMeanData_Net::* type;
    friend type get(MdHaltTag);
};
template struct Rob<MdHaltTag, &MSMeanData_Net::myHaltSpeed>;
struct MdHaltRelTag {
    typedef const double MSMeanData_Net::* type;
    friend type get(MdHaltRelTag);
};
template struct Rob<MdHaltRelTag, &MSMeanData_Net::myHaltSpeedRel>;
struct MdTrackTag {
    typedef const bool MSMeanData::* type;
    friend type get(MdTrackTag);
};
template struct Rob<MdTrackTag, &MSMeanData::myTrackVehicles>;
struct MdInternalTag {
    typedef const bool MSMeanData::* type;
    friend type get(MdInternalTag);
};
template struct Rob<MdInternalTag, &MSMeanData::myDumpInternal>;

struct NameTable {
    std::vector<int32_t> begin{0};
    std::vector<uint8_t> bytes;
    void add(const std::string& s) {
        bytes.insert(bytes.end(), s.begin(), s.end());
        begin.push_back((int32_t)bytes.size());
    }
    void put(sb2::FlatFile& ff, const std::string& prefix) {
        ff.put(prefix + "_name_begin", begin);
        ff.put(prefix + "_names", bytes);
    }
};

int edgeFlags(const MSEdge* e) {
    return (e->isInternal() ? 1 : 0) | (e->isRoundabout() ? 2 : 0) | (e->isFringe() ? 4 : 0) | (e->hasLaneChanger() ? 8 : 0)
           | (e->isCrossing() || e->isWalkingArea() ? 16 : 0);
}

// Parity harness only (sumo_flatten --trajectories-only, used by tools_dev/texttest_sweep.py): features that OBSERVE the vehicles without
// ever changing lane, position or speed -- output devices, detectors and mean data the engine has no counterpart for -- are left
// out of the flat file instead of refusing the scenario; only the trajectories are compared then.  Never set by the hook: a run
// that is asked for such an output and cannot write it is refused.
bool gTrajectoriesOnly = false;

void flattenTo(MSNet* net, sb2::FlatFile& ff) {
    OptionsCont& oc = OptionsCont::getOptions();
    // ---------------- edges and lanes
    const MSEdgeVector& edges = MSEdge::getAllEdges();
    std::map<const MSLane*, int> laneIdx;
    std::map<const MSEdge*, int> edgeIdx;
    std::map<const MSJunction*, int> junctionIdx;
    std::vector<const MSLane*> lanes;
    std::vector<const MSJunction*> junctions;
    std::vector<int32_t> edgeLaneBegin{0}, edgeFlagsV;
    NameTable edgeNames, laneNames, junctionNames;
    auto jIdx = [&](const MSJunction* j) -> int {
        if (j == nullptr) {
            return -1;
        }
        auto it = junctionIdx.find(j);
        if (it != junctionIdx.end()) {
            return it->second;
        }
        const int idx = (int)junctions.size();
        junctionIdx[j] = idx;
        junctions.push_back(j);
        junctionNames.add(j->getID());
        return idx;
    };
    for (const MSEdge* e : edges) {
        edgeIdx[e] = (int)edgeIdx.size();
        edgeNames.add(e->getID());
        edgeFlagsV.push_back(edgeFlags(e));
        for (const MSLane* l : e->getLanes()) {
            laneIdx[l] = (int)lanes.size();
            lanes.push_back(l);
            laneNames.add(l->getID());
        }
        edgeLaneBegin.push_back((int32_t)lanes.size());
    }
    auto lIdx = [&](const MSLane* l) -> int {
        if (l == nullptr) {
            return -1;
        }
        auto it = laneIdx.find(l);
        return it == laneIdx.end() ? -1 : it->second;
    };
    const int nL = (int)lanes.size();
    // ---------------- links
    std::map<const MSLink*, int> linkIdx;
    std::vector<const MSLink*> links;
    std::vector<int32_t> laneLinkBegin{0};
    for (const MSLane* l : lanes) {
        for (const MSLink* k : l->getLinkCont()) {
            linkIdx[k] = (int)links.size();
            links.push_back(k);
        }
        laneLinkBegin.push_back((int32_t)links.size());
    }
    auto kIdx = [&](const MSLink* k) -> int {
        if (k == nullptr) {
            return -1;
        }
        auto it = linkIdx.find(k);
        return it == linkIdx.end() ? -1 : it->second;
    };
    {
        std::vector<double> len, spd, wid, geomf, shapeXY, incLen;
        std::vector<int32_t> ledge, lflags, lnumid, predLog, predNorm, entryLink, incLink, incBegin{0}, incLane, incViaLink, toJ, fromJ, shapeBegin{0};
        for (const MSLane* l : lanes) {
            len.push_back(l->getLength());
            spd.push_back(l->getSpeedLimit());
            wid.push_back(l->getWidth());
            geomf.push_back(l->getLengthGeometryFactor());
            ledge.push_back(edgeIdx[&l->getEdge()]);
            lflags.push_back(edgeFlags(&l->getEdge()));
            lnumid.push_back(l->getNumericalID());
            predLog.push_back(lIdx(l->getLogicalPredecessorLane()));
            predNorm.push_back(lIdx(l->getNormalPredecessorLane()));
            entryLink.push_back(l->isInternal() ? kIdx(l->getEntryLink()) : -1);
            incLink.push_back(l->isInternal() && l->getIncomingLanes().size() == 1 ? kIdx(l->getIncomingLanes()[0].viaLink) : -1);
            for (const MSLane::IncomingLaneInfo& ili : l->getIncomingLanes()) {
                incLane.push_back(lIdx(ili.lane));
                incViaLink.push_back(kIdx(ili.viaLink));
                incLen.push_back(ili.length);
            }
            incBegin.push_back((int32_t)incLane.size());
            toJ.push_back(jIdx(l->getEdge().getToJunction()));
            fromJ.push_back(jIdx(l->getEdge().getFromJunction()));
            for (const Position& p : l->getShape()) {
                shapeXY.push_back(p.x());
                shapeXY.push_back(p.y());
            }
            shapeBegin.push_back((int32_t)(shapeXY.size() / 2));
        }
        ff.put("lane_length", len);
        ff.put("lane_speed", spd);
        ff.put("lane_width", wid);
        ff.put("lane_geomfactor", geomf);
        ff.put("lane_edge", ledge);
        ff.put("lane_flags", lflags);
        ff.put("lane_numid", lnumid);
        ff.put("lane_link_begin", laneLinkBegin);
        ff.put("lane_pred_logical", predLog);
        ff.put("lane_pred_normal", predNorm);
        ff.put("lane_entry_link", entryLink);
        ff.put("lane_inc_link", incLink);
        ff.put("lane_inc_begin", incBegin);
        ff.put("inc_lane", incLane);
        ff.put("inc_link", incViaLink);
        ff.put("inc_length", incLen);
        ff.put("lane_to_junction", toJ);
        ff.put("lane_from_junction", fromJ);
        ff.put("lane_shape_begin", shapeBegin);
        ff.put("shape_xy", shapeXY);
        laneNames.put(ff, "lane");
        ff.put("edge_lane_begin", edgeLaneBegin);
        ff.put("edge_flags", edgeFlagsV);
        edgeNames.put(ff, "edge");
    }
    // ---------------- traffic lights (index table first; links refer to it)
    std::map<const MSTrafficLightLogic*, int> tlIdx;
    std::vector<MSTrafficLightLogic*> tls;
    for (const std::string& id : net->getTLSControl().getAllTLIds()) {
        MSTrafficLightLogic* logic = net->getTLSControl().getActive(id);
        tlIdx[logic] = (int)tls.size();
        tls.push_back(logic);
    }
    {
        std::vector<int32_t> from, to, via, flags, index, tl, tlindex, intBefore, junction, entry, exitL;
        std::vector<uint8_t> state, dir;
        std::vector<double> length, foevis, radius, lbc, csize;
        std::vector<int32_t> foeLinkBegin{0}, foeLink, foeLaneBegin{0}, foeLane, cflag, cfoeidx;
        std::vector<uint64_t> response;
        for (const MSLink* k : links) {
            from.push_back(lIdx(k->getLaneBefore()));
            to.push_back(lIdx(k->getLane()));
            via.push_back(lIdx(k->getViaLane()));
            state.push_back((uint8_t)k->getState());
            dir.push_back((uint8_t)toString(k->getDirection())[0]);
            length.push_back(k->getLength());
            foevis.push_back(k->getFoeVisibilityDistance());
            radius.push_back(k->*get(RadiusTag()));
            int f = 0;
            f |= k->keepClear() ? 1 : 0;
            f |= k->hasFoes() ? 2 : 0;
            f |= k->*get(AmContTag()) ? 4 : 0;
            f |= k->isEntryLink() ? 8 : 0;
            f |= k->isConflictEntryLink() ? 16 : 0;
            f |= k->isExitLink() ? 32 : 0;
            f |= k->isExitLinkAfterInternalJunction() ? 64 : 0;
            f |= k->isInternalJunctionLink() ? 128 : 0;
            f |= k->isIndirect() ? 256 : 0;
            f |= k->*get(PedFoeTag()) ? 512 : 0;
            f |= k->isTurnaround() ? 1024 : 0;
            f |= k->*get(AmContOffTag()) ? 2048 : 0;
            flags.push_back(f);
            index.push_back(k->getIndex());
            auto ti = tlIdx.find(k->getTLLogic());
            tl.push_back(ti == tlIdx.end() ? -1 : ti->second);
            tlindex.push_back(k->getTLIndex());
            intBefore.push_back(lIdx(k->getInternalLaneBefore()));
            junction.push_back(jIdx(k->getJunction()));
            entry.push_back(kIdx(k->getCorrespondingEntryLink()));
            exitL.push_back(kIdx(k->getCorrespondingExitLink()));
            for (const MSLink* fl : k->getFoeLinks()) {
                foeLink.push_back(kIdx(fl));
            }
            foeLinkBegin.push_back((int32_t)foeLink.size());
            const std::vector<const MSLane*>& fls = k->getFoeLanes();
            const std::vector<MSLink::ConflictInfo>& cis = k->getConflicts();
            // (MSLink::recheckSetRequestInformation, MSLink.cpp:614, appends conflict records WITHOUT a foe lane: they are only reached
            // through the foeConflictIndex of the other link's record -- exported with foe lane -1)
            for (int i = 0; i < (int)std::max(fls.size(), cis.size()); i++) {
                foeLane.push_back(i < (int)fls.size() ? lIdx(fls[i]) : -1);
                if (i < (int)cis.size()) {
                    lbc.push_back(cis[i].lengthBehindCrossing);
                    csize.push_back(cis[i].conflictSize);
                    cflag.push_back((int)cis[i].flag);
                    cfoeidx.push_back(cis[i].foeConflictIndex);
                } else {
                    // entry links keep foe lanes without conflict geometry (MSLink.cpp:249: geometry only if internalLaneBefore)
                    lbc.push_back(0);
                    csize.push_back(0);
                    cflag.push_back(-1);
                    cfoeidx.push_back(-1);
                }
            }
            foeLaneBegin.push_back((int32_t)foeLane.size());
            uint64_t mask[4] = {0, 0, 0, 0};
            const MSJunction* j = k->getJunction();
            if (j != nullptr && j->getLogic() != nullptr && k->getIndex() >= 0 && k->getIndex() < j->getLogic()->getLogicSize()) {
                const MSLogicJunction::LinkBits& bits = j->getLogic()->getResponseFor(k->getIndex());
                for (int b = 0; b < 256 && b < (int)bits.size(); b++) {
                    if (bits.test(b)) {
                        mask[b / 64] |= (uint64_t)1 << (b % 64);
                    }
                }
            }
            for (int w = 0; w < 4; w++) {
                response.push_back(mask[w]);
            }
        }
        ff.put("link_from", from);
        ff.put("link_to", to);
        ff.put("link_via", via);
        ff.put("link_state", state);
        ff.put("link_dir", dir);
        ff.put("link_length", length);
        ff.put("link_foevis", foevis);
        ff.put("link_radius", radius);
        ff.put("link_flags", flags);
        ff.put("link_index", index);
        ff.put("link_tl", tl);
        ff.put("link_tlindex", tlindex);
        ff.put("link_internal_before", intBefore);
        ff.put("link_junction", junction);
        ff.put("link_entry", entry);
        ff.put("link_exit", exitL);
        ff.put("link_foelink_begin", foeLinkBegin);
        ff.put("foelink", foeLink);
        ff.put("link_foelane_begin", foeLaneBegin);
        ff.put("foelane", foeLane);
        ff.put("conf_lbc", lbc);
        ff.put("conf_size", csize);
        ff.put("conf_flag", cflag);
        ff.put("conf_foeidx", cfoeidx);
        ff.put("link_response", response);
        junctionNames.put(ff, "junction");
    }
    // ---------------- static traffic light programs
    {
        std::vector<int32_t> phaseBegin{0}, stateBegin{0}, curPhase, tlLinkBegin{0}, tlLinkIndex, tlLinkLink, tlStatic;
        std::vector<int64_t> dur, nextSwitch;
        std::vector<uint8_t> states;
        NameTable tlNames;
        for (MSTrafficLightLogic* logic : tls) {
            tlNames.add(logic->getID());
            if (dynamic_cast<MSSimpleTrafficLightLogic*>(logic) == nullptr || logic->getLogicType() != TrafficLightType::STATIC) {
                throw ProcessError("sumob200: traffic light '" + logic->getID() + "' is of type '" + toString(logic->getLogicType()) + "' (only static programs are supported)");
            }
            tlStatic.push_back(dynamic_cast<MSSimpleTrafficLightLogic*>(logic) != nullptr && logic->getLogicType() == TrafficLightType::STATIC ? 1 : 0);
            for (const MSPhaseDefinition* p : logic->getPhases()) {
                dur.push_back(p->duration);
                states.insert(states.end(), p->getState().begin(), p->getState().end());
                stateBegin.push_back((int32_t)states.size());
            }
            phaseBegin.push_back((int32_t)dur.size());
            curPhase.push_back(logic->getCurrentPhaseIndex());
            nextSwitch.push_back(logic->getNextSwitchTime());
            const MSTrafficLightLogic::LinkVectorVector& lvv = logic->getLinks();
            for (int i = 0; i < (int)lvv.size(); i++) {
                for (const MSLink* k : lvv[i]) {
                    tlLinkIndex.push_back(i);
                    tlLinkLink.push_back(kIdx(k));
                }
            }
            tlLinkBegin.push_back((int32_t)tlLinkIndex.size());
        }
        ff.put("tl_phase_begin", phaseBegin);
        ff.put("phase_dur", dur);
        ff.put("phase_state_begin", stateBegin);
        ff.put("phase_states", states);
        ff.put("tl_cur_phase", curPhase);
        ff.put("tl_next_switch", nextSwitch);
        ff.put("tl_link_begin", tlLinkBegin);
        ff.put("tllink_index", tlLinkIndex);
        ff.put("tllink_link", tlLinkLink);
        ff.put("tl_static", tlStatic);
        tlNames.put(ff, "tl");
    }
    // ---------------- vehicles, their types and routes (everything loaded: run with --route-steps 0)
    {
        MSVehicleControl& vc = net->getVehicleControl();
        std::vector<const SUMOVehicle*> vehs;
        for (auto it = vc.loadedVehBegin(); it != vc.loadedVehEnd(); ++it) {
            vehs.push_back(it->second);
        }
        // insertion order: MSVehicleContainer keys by depart time, equal times keep load order (= numerical id order)
        // (MSVehicleContainer::add buckets by depart rounded UP to a step boundary, MSVehicleContainer.cpp:73-89)
        auto bucket = [](const SUMOVehicle* a) {
            const SUMOTime d = a->getParameter().depart;
            const SUMOTime delay = d % DELTA_T;
            return d + (delay == 0 ? 0 : DELTA_T - delay);
        };
        std::stable_sort(vehs.begin(), vehs.end(), [&](const SUMOVehicle* a, const SUMOVehicle* b) {
            if (bucket(a) != bucket(b)) {
                return bucket(a) < bucket(b);
            }
            return a->getNumericalID() < b->getNumericalID();
        });
        // ---- features of the demand the step engine does not model: refused here, never silently dropped
        if (net->getInsertionControl().getPendingFlowCount() > 0) {
            throw ProcessError("sumob200: <flow> definitions are not supported (expand them into <vehicle>s)");
        }
        if (net->hasPersons() || net->hasContainers()) {
            throw ProcessError("sumob200: persons / containers are not supported");
        }
        for (const SUMOVehicle* v : vehs) {
            const SUMOVehicleParameter& p = v->getParameter();
            const std::string who = "sumob200: vehicle '" + v->getID() + "': ";
            if (!p.stops.empty() || !v->getRoute().getStops().empty() || v->hasStops()) {
                throw ProcessError(who + "<stop>s are not supported");
            }
            if (p.arrivalSpeedProcedure != ArrivalSpeedDefinition::DEFAULT && p.arrivalSpeedProcedure != ArrivalSpeedDefinition::CURRENT
                    && p.arrivalSpeedProcedure != ArrivalSpeedDefinition::GIVEN) {
                throw ProcessError(who + "arrivalSpeed must be 'current' or a number");
            }
            // (edges have one lane: a given or first-allowed arrival lane is lane 0, where the vehicle is anyway; MSVehicle::updateBestLanes
            // then treats the arrival like a stop, which only shortens the LaneQ lengths lane changing would use.  random draws.)
            if (p.arrivalLaneProcedure == ArrivalLaneDefinition::RANDOM) {
                throw ProcessError(who + "arrivalLane is not supported");
            }
            if (p.departEdgeProcedure != RouteIndexDefinition::DEFAULT || p.arrivalEdgeProcedure != RouteIndexDefinition::DEFAULT) {
                throw ProcessError(who + "departEdge / arrivalEdge are not supported");
            }
            if (p.insertionChecks != (int)InsertionCheck::ALL) {
                throw ProcessError(who + "insertionChecks other than 'all' are not supported");
            }
            const SUMOVTypeParameter& tp = v->getVehicleType().getParameter();
            if (tp.actionStepLength != 0 && tp.actionStepLength != DELTA_T) {
                throw ProcessError(who + "actionStepLength different from the step length is not supported");
            }
            if (!tp.jmParameter.empty()) {
                throw ProcessError(who + "junction model parameters (jm*) are not supported");
            }
            for (const MSVehicleDevice* d : v->getDevices()) {
                const std::string dev = d->deviceName();
                const bool observer = dev == "ssm" || dev == "btreceiver" || dev == "btsender" || dev == "battery" || dev == "example";
                if (dev != "emissions" && dev != "tripinfo" && dev != "fcd" && dev != "vehroute" && !(gTrajectoriesOnly && observer)) {
                    throw ProcessError(who + "device '" + dev + "' is not supported");
                }
            }
            for (const MSEdge* e : v->getRoute().getEdges()) {
                for (const MSLane* l : e->getLanes()) {
                    if (!l->allowsVehicleClass(v->getVClass())) {
                        throw ProcessError(who + "lane '" + l->getID() + "' on its route does not allow its vehicle class (lane permissions are not modelled)");
                    }
                }
            }
        }
        std::map<const MSVehicleType*, int> typeIdx;
        std::vector<const MSVehicleType*> types;
        std::map<const MSRoute*, int> routeIdx;
        std::vector<const MSRoute*> routes;
        std::vector<int32_t> numid, vtype, vroute, dpp, dsp, dlp, dlane, idrank(vehs.size(), 0);
        std::vector<uint8_t> hasEmissions;   // vehicle carries MSDevice_Emissions (--device.emissions.probability / explicit)
        std::vector<int64_t> depart;
        std::vector<double> sf, arrpos, arrspeed, dpos, dspeed;
        NameTable vehNames;
        for (const SUMOVehicle* v : vehs) {
            const MSVehicleType* t = &v->getVehicleType();
            if (typeIdx.count(t) == 0) {
                typeIdx[t] = (int)types.size();
                types.push_back(t);
            }
            const MSRoute* r = &v->getRoute();
            if (routeIdx.count(r) == 0) {
                routeIdx[r] = (int)routes.size();
                routes.push_back(r);
            }
            const SUMOVehicleParameter& p = v->getParameter();
            numid.push_back((int32_t)v->getNumericalID());
            vtype.push_back(typeIdx[t]);
            vroute.push_back(routeIdx[r]);
            depart.push_back(p.depart);
            sf.push_back(v->getChosenSpeedFactor());
            arrpos.push_back(v->getArrivalPos());
            arrspeed.push_back(p.arrivalSpeedProcedure == ArrivalSpeedDefinition::GIVEN ? p.arrivalSpeed : -1.);
            dpp.push_back((int)p.departPosProcedure);
            dpos.push_back(p.departPos);
            dsp.push_back((int)p.departSpeedProcedure);
            dspeed.push_back(p.departSpeed);
            dlp.push_back((int)p.departLaneProcedure);
            dlane.push_back(p.departLane);
            hasEmissions.push_back(v->getDevice(typeid(MSDevice_Emissions)) != nullptr ? 1 : 0);
            vehNames.add(v->getID());
        }
        {
            // rank in std::string order (MSVehicle::isLeader tie breaker MSVehicle.cpp:7791; FCD output order MSFCDExport.cpp:98)
            std::vector<int> order(vehs.size());
            for (int i = 0; i < (int)order.size(); i++) {
                order[i] = i;
            }
            std::sort(order.begin(), order.end(), [&](int a, int b) {
                return vehs[a]->getID() < vehs[b]->getID();
            });
            for (int r = 0; r < (int)order.size(); r++) {
                idrank[order[r]] = r;
            }
        }
        ff.put("veh_numid", numid);
        ff.put("veh_type", vtype);
        ff.put("veh_route", vroute);
        ff.put("veh_depart", depart);
        ff.put("veh_speedfactor", sf);
        ff.put("veh_arrivalpos", arrpos);
        ff.put("veh_arrivalspeed", arrspeed);
        ff.put("veh_departpos_proc", dpp);
        ff.put("veh_departpos", dpos);
        ff.put("veh_departspeed_proc", dsp);
        ff.put("veh_departspeed", dspeed);
        ff.put("veh_departlane_proc", dlp);
        ff.put("veh_departlane", dlane);
        ff.put("veh_idrank", idrank);
        ff.put("veh_emissions", hasEmissions);
        vehNames.put(ff, "veh");
        // types
        std::vector<double> tl, tg, tms, tdms, ta, td, ted, tad, ttau, tsig, timp, tw, tcmg, idmDelta, idmStep;
        std::vector<int32_t> tcf, tem, temModel;
        std::vector<double> temCoef, temScale, temMass, temIncl;
        std::vector<int64_t> tsd;
        NameTable typeNames;
        for (const MSVehicleType* t : types) {
            const MSCFModel& cf = t->getCarFollowModel();
            typeNames.add(t->getID());
            tl.push_back(t->getLength());
            tg.push_back(t->getMinGap());
            tms.push_back(t->getMaxSpeed());
            tdms.push_back(t->getDesiredMaxSpeed());
            ta.push_back(cf.getMaxAccel());
            td.push_back(cf.getMaxDecel());
            ted.push_back(cf.getEmergencyDecel());
            tad.push_back(cf.getApparentDecel());
            ttau.push_back(cf.getHeadwayTime());
            tsig.push_back(cf.getImperfection());
            timp.push_back(t->getImpatience());
            tw.push_back(t->getWidth());
            tcmg.push_back(cf.getCollisionMinGapFactor());
            tcf.push_back(cf.getModelID());
            tsd.push_back(cf.getStartupDelay());
            tem.push_back((int)t->getEmissionClass());
            {
                // emission model of the type: 0 = none/unsupported (zero emissions are NOT assumed: the engine refuses the
                // emission device for such a type), 3 = HelpersHBEFA3::compute, 4 = HelpersHBEFA4::compute (gasoline/diesel)
                const SUMOEmissionClass ec = t->getEmissionClass();
                const PollutantsInterface::Helper* h = PollutantsInterface::myHelpers[ec >> 16];
                // PollutantsInterface::Helper::myVolumetricFuel is set from this option (PollutantsInterface.cpp:116)
                const bool volumetricFuel = OptionsCont::getOptions().exists("emissions.volumetric-fuel") && OptionsCont::getOptions().getBool("emissions.volumetric-fuel");
                int model = 0;
                double coef[7][7] = {{0}};
                double scale[7] = {1, 1, 1, 1, 1, 1, 1};
                if (h == &PollutantsInterface::myHBEFA3Helper) {
                    model = 3;
                    const int index = (ec & ~PollutantsInterface::HEAVY_BIT) - HelpersHBEFA3::HBEFA3_BASE;
                    for (int e = 0; e < 6; e++) {
                        for (int k = 0; k < 6; k++) {
                            coef[e][k] = HelpersHBEFA3::myFunctionParameter[index][e][k];
                        }
                        scale[e] = 3.6;
                        if (e == PollutantsInterface::FUEL && volumetricFuel) {
                            scale[e] *= (h->getFuel(ec) == "Diesel") ? 836. : 742.;
                        }
                    }
                } else if (h == &PollutantsInterface::myHBEFA4Helper && h->getFuel(ec) != "Electricity") {
                    model = 4;
                    const int index = (ec & ~PollutantsInterface::HEAVY_BIT) - HelpersHBEFA4::HBEFA4_BASE;
                    for (int e = 0; e < 7; e++) {
                        for (int k = 0; k < 7; k++) {
                            coef[e][k] = HelpersHBEFA4::myFunctionParameter[index][e][k];
                        }
                        if (e == PollutantsInterface::FUEL && volumetricFuel) {
                            if (h->getFuel(ec) == "Diesel") {
                                scale[e] *= 836.;
                            } else if (h->getFuel(ec) == "Gasoline") {
                                scale[e] *= 742.;
                            }
                        }
                    }
                }
                temModel.push_back(model);
                for (int e = 0; e < 7; e++) {
                    for (int k = 0; k < 7; k++) {
                        temCoef.push_back(coef[e][k]);
                    }
                    temScale.push_back(scale[e]);
                }
                // PollutantsInterface::Helper::getCoastingDecel (PollutantsInterface.cpp:197-210): mass and the speed coefficient
                const EnergyParams* ep = t->getEmissionParameters();
                const double mass = ep->getDouble(SUMO_ATTR_MASS);
                temMass.push_back(mass);
                temIncl.push_back(ep->getDouble(SUMO_ATTR_FRONTSURFACEAREA) / mass * -9.05337017 + -0.00017774);
            }
            if (TIME2STEPS(t->getParameter().getCFParam(SUMO_ATTR_SIGMA_STEP, TS)) > DELTA_T) {
                // MSCFModel_Krauss::myDawdleStep: the dawdle draw is kept for several steps (MSCFModel_Krauss.cpp:51-60, 96-113)
                throw ProcessError("sumob200: vType '" + t->getID() + "': sigmaStep longer than the step length is not supported");
            }
            if (!t->getCarFollowModel().getMaxAccelProfile().empty() || !t->getCarFollowModel().getDesAccelProfile().empty()) {
                throw ProcessError("sumob200: vType '" + t->getID() + "': maxAccelProfile / desAccelProfile are not supported");
            }
            if (t->getParameter().getTimeToTeleport(-12345) != -12345 || t->getParameter().getTimeToTeleportBidi(-12345) != -12345) {
                throw ProcessError("sumob200: vType '" + t->getID() + "': timeToTeleport per vType is not supported");
            }
            idmDelta.push_back(t->getParameter().getCFParam(SUMO_ATTR_CF_IDM_DELTA, 4.));
            idmStep.push_back(t->getParameter().getCFParam(SUMO_ATTR_CF_IDM_STEPPING, .25));
        }
        ff.put("vt_length", tl);
        ff.put("vt_mingap", tg);
        ff.put("vt_maxspeed", tms);
        ff.put("vt_desiredmaxspeed", tdms);
        ff.put("vt_accel", ta);
        ff.put("vt_decel", td);
        ff.put("vt_emergencydecel", ted);
        ff.put("vt_apparentdecel", tad);
        ff.put("vt_tau", ttau);
        ff.put("vt_sigma", tsig);
        ff.put("vt_impatience", timp);
        ff.put("vt_width", tw);
        ff.put("vt_collision_mingap_factor", tcmg);
        ff.put("vt_cfmodel", tcf);
        ff.put("vt_startupdelay", tsd);
        ff.put("vt_emission_class", tem);
        ff.put("vt_em_model", temModel);
        ff.put("vt_em_coef", temCoef);
        ff.put("vt_em_scale", temScale);
        ff.put("vt_em_mass", temMass);
        ff.put("vt_em_incl", temIncl);
        ff.put("vt_idm_delta", idmDelta);
        ff.put("vt_idm_stepping", idmStep);
        typeNames.put(ff, "vt");
        // routes
        std::vector<int32_t> routeBegin{0}, routeEdges;
        for (const MSRoute* r : routes) {
            for (const MSEdge* e : r->getEdges()) {
                routeEdges.push_back(edgeIdx[e]);
            }
            routeBegin.push_back((int32_t)routeEdges.size());
        }
        ff.put("route_begin", routeBegin);
        ff.put("route_edges", routeEdges);
    }
    // ---------------- other things the step engine does not model
    // ---- options: everything of the "Processing" and "Mesoscopic" topics that is not at its default must be one the engine is known to
    // honour (exported below, or applied by the reference's loader to the objects that are exported) -- never silently ignored
    {
        static const std::set<std::string> honoured = {
            "route-steps", "scale", "scale-suffix",  // loader only (the hook loads everything at once; --scale has been applied to the loaded vehicles)
            "default.speeddev", "default.departspeed", "default.departlane", "default.carfollowmodel", "default.emergencydecel", "extrapolate-departpos",   // in the exported vTypes / scalars
            "collision.mingap-factor",               // MSVehicleType::getCollisionMinGapFactor, exported per vType
            "time-to-teleport", "time-to-impatience", "ignore-junction-blocker", "startup-wait-threshold", "tls.yellow.min-decel",
            "emergencydecel.warning-threshold",      // exported scalars
            "step-method.ballistic", "lateral-resolution", "no-internal-links", "lanechange.duration", "ignore-route-errors", "max-depart-delay",   // refused with their own message
            "waiting-time-memory", "lanechange.overtake-right", "pedestrian.model", "keep-after-arrival", "insertion-checks",
            "tls.actuated.show-detectors", "tls.actuated.detector-length", "tls.actuated.jam-threshold", "parking.maneuver",
            "weights.reversal-penalty", "weights.turnaround-penalty", "mapmatch.distance", "mapmatch.junctions", "mapmatch.taz", "slope-centered",
            "persontrip.walk-opposite-factor", "ride.stop-tolerance", "railsignal-moving-block", "railsignal.max-block-length",
            "railsignal.default-classes", "railsignal.moving-block.default-classes", "railsignal.moving-block.max-dist",
            // (lane changing, persons, rail, actuated lights, stops, routing: the objects they act on are refused)
        };
        for (const std::string& topic : {std::string("Processing"), std::string("Mesoscopic")}) {
            for (const std::string& name : oc.getSubTopicsEntries(topic)) {
                if (oc.isSet(name) && !oc.isDefault(name) && honoured.count(name) == 0 && name.compare(0, 11, "pedestrian.") != 0
                        && name.compare(0, 14, "overhead-wire.") != 0) {
                    throw ProcessError("sumob200: option --" + name + " is not modelled by the engine (only its default is supported)");
                }
            }
        }
    }
    if (oc.isSet("load-state")) {
        throw ProcessError("sumob200: --load-state is not supported");
    }
    if (oc.getBool("ignore-route-errors")) {
        throw ProcessError("sumob200: --ignore-route-errors is not supported (insertion with impossible depart speeds)");
    }
    if (string2time(oc.getString("max-depart-delay")) >= 0 && !gTrajectoriesOnly) {
        // (the engine discards such vehicles, opt_max_depart_delay below; the hook does not delete them from MSVehicleControl yet)
        throw ProcessError("sumob200: --max-depart-delay is not supported");
    }
    for (const MSEdge* e : MSEdge::getAllEdges()) {
        for (const MSLane* l : e->getLanes()) {
            if (l->getBidiLane() != nullptr) {
                throw ProcessError("sumob200: bidirectional lanes (bidi) are not supported");
            }
            if (l->getOpposite() != nullptr) {
                throw ProcessError("sumob200: opposite-direction driving (<neigh> lanes) is not supported");
            }
        }
    }
    // ---------------- triggers of additional files change lane speeds, routes or flows during the run: not modelled, refused
    if (!MSLaneSpeedTrigger::getInstances().empty()) {
        throw ProcessError("sumob200: variableSpeedSigns are not supported");
    }
    if (!MSTriggeredRerouter::getInstances().empty()) {
        throw ProcessError("sumob200: rerouters are not supported");
    }
    if (!MSCalibrator::getInstances().empty()) {
        throw ProcessError("sumob200: calibrators are not supported");
    }
    // ---------------- E1 induction loops (NLDetectorBuilder::buildInductLoop; row f4): lane, position, length, in id order (the
    // order MSDetectorControl::writeOutput visits them).  Other detector types are not modelled by the engine: refused.
    {
        MSDetectorControl& dc = net->getDetectorControl();
        for (const SumoXMLTag tag : dc.getAvailableTypes()) {
            if (tag != SUMO_TAG_INDUCTION_LOOP && tag != SUMO_TAG_LANE_AREA_DETECTOR && dc.getTypedDetectors(tag).size() > 0 && !gTrajectoriesOnly) {
                throw ProcessError("sumob200: detectors of type '" + toString(tag) + "' are not supported (only inductionLoop / e1Detector and laneAreaDetector / e2Detector)");
            }
        }
        std::vector<int32_t> e1Lane;
        std::vector<double> e1Pos, e1Length;
        NameTable e1Names;
        for (const auto& item : dc.getTypedDetectors(SUMO_TAG_INDUCTION_LOOP)) {
            const MSInductLoop* loop = dynamic_cast<const MSInductLoop*>(item.second);
            if (loop == nullptr) {
                throw ProcessError("sumob200: unexpected detector object '" + item.first + "'");
            }
            if (gTrajectoriesOnly && (!loop->getVehicleTypes().empty() || loop->detectsPersons())) {
                continue;
            }
            if (!loop->getVehicleTypes().empty() || loop->detectsPersons()) {
                throw ProcessError("sumob200: detector '" + item.first + "': vTypes / detectPersons filters are not supported");
            }
            e1Lane.push_back(lIdx(loop->getLane()));
            e1Pos.push_back(loop->getPosition());
            e1Length.push_back(loop->getEndPosition() - loop->getPosition());
            e1Names.add(item.first);
        }
        if (!e1Lane.empty()) {
            ff.put("e1_lane", e1Lane);
            ff.put("e1_pos", e1Pos);
            ff.put("e1_length", e1Length);
            e1Names.put(ff, "e1");
        }
        // ---- E2 lane area detectors (NLDetectorBuilder::buildE2Detector; row f4): the built detector's lane, start, length and jam
        // thresholds, in id order.  Refused: detectors over several lanes, filters, detectors coupled to a traffic light (they write
        // when the light switches, not periodically).
        std::set<const MSDetectorFileOutput*> periodic;
        for (const auto& iv : dc.myIntervals) {
            for (const auto& df : iv.second) {
                periodic.insert(df.first);
            }
        }
        std::vector<int32_t> e2Lane;
        std::vector<double> e2Start, e2Length, e2JamSpeed, e2JamDist;
        std::vector<int64_t> e2JamTime;
        NameTable e2Names;
        for (const auto& item : dc.getTypedDetectors(SUMO_TAG_LANE_AREA_DETECTOR)) {
            MSE2Collector* det = dynamic_cast<MSE2Collector*>(item.second);
            if (det == nullptr) {
                throw ProcessError("sumob200: unexpected detector object '" + item.first + "'");
            }
            const std::string who = "sumob200: detector '" + item.first + "': ";
            if (gTrajectoriesOnly && (det->myLanes.size() != 1 || !det->getVehicleTypes().empty() || det->detectsPersons() || !det->myNextEdges.empty()
                                      || periodic.count(det) == 0)) {
                continue;
            }
            if (det->myLanes.size() != 1) {
                throw ProcessError(who + "lane area detectors over several lanes are not supported");
            }
            if (!det->getVehicleTypes().empty() || det->detectsPersons() || !det->myNextEdges.empty()) {
                throw ProcessError(who + "vTypes / nextEdges / detectPersons filters are not supported");
            }
            if (periodic.count(det) == 0) {
                throw ProcessError(who + "lane area detectors coupled to a traffic light are not supported");
            }
            e2Lane.push_back(lIdx(det->getLane()));
            e2Start.push_back(det->getStartPos());
            e2Length.push_back(det->getLength());
            e2JamSpeed.push_back(det->myJamHaltingSpeedThreshold);
            e2JamDist.push_back(det->myJamDistanceThreshold);
            e2JamTime.push_back((int64_t)det->myJamHaltingTimeThreshold);
            e2Names.add(item.first);
        }
        if (!e2Lane.empty()) {
            ff.put("e2_lane", e2Lane);
            ff.put("e2_start", e2Start);
            ff.put("e2_length", e2Length);
            ff.put("e2_jam_speed", e2JamSpeed);
            ff.put("e2_jam_dist", e2JamDist);
            ff.put("e2_jam_time", e2JamTime);
            e2Names.put(ff, "e2");
        }
    }
    // ---------------- edgeData / laneData (MSMeanData_Net; row f4): the engine accumulates the raw per-lane members, the writer stays here
    {
        bool any = false;
        double halt = 0;
        for (const auto& item : net->getDetectorControl().getMeanData()) {
            for (const MSMeanData* md : item.second) {
                const MSMeanData_Net* mdn = dynamic_cast<const MSMeanData_Net*>(md);
                if (gTrajectoriesOnly && (mdn == nullptr || md->*get(MdTrackTag()) || md->*get(MdInternalTag()) || mdn->*get(MdHaltRelTag()) > 0
                                          || !md->getVehicleTypes().empty() || (any && mdn->*get(MdHaltTag()) != halt))) {
                    continue;
                }
                if (mdn == nullptr) {
                    throw ProcessError("sumob200: mean data '" + item.first + "' is not of type edgeData / laneData (emissions, harmonoise, ... are not supported)");
                }
                if (md->*get(MdTrackTag()) || md->*get(MdInternalTag()) || mdn->*get(MdHaltRelTag()) > 0 || !md->getVehicleTypes().empty()) {
                    throw ProcessError("sumob200: mean data '" + item.first + "': trackVehicles, withInternal, speedThresholdRelative and vTypes are not supported");
                }
                const double hs = mdn->*get(MdHaltTag());
                if (any && hs != halt) {
                    throw ProcessError("sumob200: mean data definitions with different speedThreshold values are not supported");
                }
                any = true;
                halt = hs;
            }
        }
        if (any) {
            ff.putScalar<int32_t>("md_enabled", 1);
            ff.putScalar<double>("md_halt_speed", halt);
        }
    }
    // ---------------- options the step depends on (MSFrame::setMSGlobals, MSFrame.cpp:1172-1283)
    ff.putScalar<int64_t>("opt_delta_t", DELTA_T);
    ff.putScalar<int64_t>("opt_begin", string2time(oc.getString("begin")));
    ff.putScalar<int64_t>("opt_end", string2time(oc.getString("end")));
    ff.putScalar<int64_t>("opt_seed", (int64_t)oc.getInt("seed"));
    ff.putScalar<int64_t>("opt_time_to_teleport", MSGlobals::gTimeToGridlock);
    ff.putScalar<int64_t>("opt_max_depart_delay", string2time(oc.getString("max-depart-delay")));
    ff.putScalar<int32_t>("opt_extrapolate_departpos", oc.getBool("extrapolate-departpos") ? 1 : 0);
    ff.putScalar<int64_t>("opt_time_to_impatience", MSGlobals::gTimeToImpatience);
    ff.putScalar<int64_t>("opt_startup_wait_threshold", MSGlobals::gStartupWaitThreshold);
    ff.putScalar<int64_t>("opt_ignore_junction_blocker", MSGlobals::gIgnoreJunctionBlocker);
    ff.putScalar<double>("opt_tls_yellow_min_decel", MSGlobals::gTLSYellowMinDecel);
    ff.putScalar<double>("opt_emergency_decel_warning_threshold", MSGlobals::gEmergencyDecelWarningThreshold);
    ff.putScalar<int32_t>("opt_default_departspeed_proc", (int)MSLane::getDefaultDepartSpeedDefinition());
    ff.putScalar<double>("opt_default_departspeed", MSLane::getDefaultDepartSpeed());
    ff.putScalar<int32_t>("opt_using_internal_lanes", MSGlobals::gUsingInternalLanes ? 1 : 0);
    ff.putScalar<int32_t>("opt_semi_implicit_euler", MSGlobals::gSemiImplicitEulerUpdate ? 1 : 0);
    ff.putScalar<int32_t>("opt_sublane", MSGlobals::gSublane ? 1 : 0);
    ff.putScalar<int32_t>("opt_thread_rngs", oc.getInt("thread-rngs"));
    ff.putScalar<double>("opt_lanechange_duration", STEPS2TIME(MSGlobals::gLaneChangeDuration));
    // fleet aggregates behind MSLane::getMaximumBrakeDist (MSLane.cpp:2945-2953)
    ff.putScalar<double>("opt_max_speed_factor", net->getVehicleControl().getMaxSpeedFactor());
    ff.putScalar<double>("opt_min_deceleration", net->getVehicleControl().getMinDeceleration());
    ff.putScalar<double>("opt_max_mingap", net->getVehicleControl().getMaxMinGap());
    // numeric values of the car-following model tags written to vt_cfmodel (MSCFModel::getModelID)
    ff.putScalar<int32_t>("opt_tag_cf_krauss", (int)SUMO_TAG_CF_KRAUSS);
    ff.putScalar<int32_t>("opt_tag_cf_krauss_orig1", (int)SUMO_TAG_CF_KRAUSS_ORIG1);
    ff.putScalar<int32_t>("opt_tag_cf_idm", (int)SUMO_TAG_CF_IDM);
    ff.putScalar<int32_t>("opt_default_departlane_proc", (int)MSEdge::getAllEdges().front()->getDefaultDepartLaneDefinition());
    ff.putScalar<int32_t>("opt_default_departlane", MSEdge::getAllEdges().front()->getDefaultDepartLane());
}

void flatten(MSNet* net, const std::string& path) {
    sb2::FlatFile ff;
    flattenTo(net, ff);
    ff.save(path);
}

// ---- <flow>s (parity harness only; the hook refuses them).  The reference turns a flow into vehicles step by step:
// MSNet::simulationStep calls MSInsertionControl::determineCandidates(time) (MSInsertionControl.cpp:219-301), which builds the
// vehicles due at `time` (ids "flow.N", type / speedFactor drawn from the route parser's RNG, departure of probabilistic flows
// from myFlowRNG), files them in MSVehicleContainer's bucket of their depart time and moves the buckets that are due to
// myPendingEmits.  The engine wants every vehicle up front, so the harness makes the reference's own code do the expansion
// BEFORE the run: the same calls, in the same order (time ascending), hence the same draws from both RNGs, the same numerical ids
// and the same order inside each depart bucket (loaded vehicles first, then the flows' in flow order).  What checkCandidates
// moved to myPendingEmits on the way is filed back (MSVehicleContainer::add appends to the bucket, the order is kept), so the run
// that follows sees ordinary pre-loaded vehicles.  tools_dev/texttest_sweep.py pins this against the untouched reference: the
// trace of a run without --flat-out (no expansion) must equal the trace of the expanded run byte by byte.
// Returns the time of the step in which the reference would have dropped its last flow (a run without --end goes on until then even
// if the road is empty), -1 without flows.
SUMOTime expandFlows(MSNet* net, SUMOTime begin, SUMOTime end) {
    MSInsertionControl& ic = net->getInsertionControl();
    if (ic.getPendingFlowCount() == 0) {
        return -1;
    }
    if (MSRoutingEngine::isEnabled()) {
        throw ProcessError("sumob200: <flow>s with the routing engine enabled are not supported");
    }
    if (!ic.myPendingEmits.empty()) {
        throw ProcessError("sumob200: <flow> expansion: insertion backlog before the first step");
    }
    const SUMOTime horizon = end >= 0 ? end : begin + 1000000 * DELTA_T;
    SUMOTime last = -1;
    for (SUMOTime t = begin; t < horizon && ic.getPendingFlowCount() > 0; t += DELTA_T) {
        ic.determineCandidates(t);
        last = t;
    }
    if (ic.getPendingFlowCount() > 0) {
        if (end < 0) {
            throw ProcessError("sumob200: <flow> without an end in a simulation without --end");
        }
        // flows that outlive --end: nothing more of them departs in this run (probabilistic ones would draw again)
        for (MSInsertionControl::Flow& f : ic.myFlows) {
            delete f.pars;
        }
        ic.myFlows.clear();
    }
    MSVehicleContainer::VehicleVector due;
    due.swap(ic.myPendingEmits);
    for (SUMOVehicle* v : due) {
        ic.myAllVeh.add(v);
    }
    return last;
}

// ---- per-vehicle dawdle draws of the coming step (sigma > 0 parity, BASELINE north_star "replaying the reference's
// per-vehicle RNG draws").  Nothing in the reference is patched: the draws are PREDICTED from public state before the step
// and the prediction is verified against the streams' public call counters after the step.
//   * MSCFModel_Krauss::patchSpeedBeforeLC draws once per executeMove from veh->getRNG() = the RNG of the vehicle's lane
//     (MSCFModel_Krauss.cpp:117,165; MSBaseVehicle::getRNG; MSLane.h:246), value = rng() / 2^32 (RandHelper.cpp:105)
//   * executeMove order = MSEdgeControl::myActiveLanes order (written by MSEdgeControl::saveState, MSEdgeControl.cpp:389),
//     within a lane front to back (MSLane.cpp:2321)
struct DrawPrediction {
    std::vector<std::pair<int64_t, double> > draws;                 // (vehicle numerical id, draw)
    std::map<SumoRNG*, std::pair<unsigned long long, int> > streams;   // stream -> (count before, predicted calls)
    void checkAndWrite(SUMOTime t, FILE* fd) {
        int64_t ok = 1;
        for (const auto& it : streams) {
            if (it.first->count != it.second.first + (unsigned long long)it.second.second) {
                ok = 0;
                if (getenv("SB2_DRAW_DEBUG") != nullptr) {
                    fprintf(stderr, "draws t=%lld stream %s: predicted %d calls, saw %lld\n", (long long)t, it.first->id.c_str(), it.second.second,
                            (long long)(it.first->count - it.second.first));
                }
            }
        }
        const int64_t hdr[3] = {(int64_t)t, (int64_t)draws.size(), ok};
        fwrite(hdr, 8, 3, fd);
        for (const auto& d : draws) {
            fwrite(&d.first, 8, 1, fd);
            fwrite(&d.second, 8, 1, fd);
        }
    }
};

DrawPrediction
predictDraws(MSNet* net) {
    DrawPrediction p;
    OutputDevice_String od(0);
    net->getEdgeControl().saveState(od);
    const std::string xml = od.getString();
    const size_t a = xml.find("lanes=\"");
    if (a == std::string::npos) {
        return p;
    }
    const size_t b = xml.find('"', a + 7);
    std::map<SumoRNG*, std::mt19937> clones;
    std::list<MSLane*> active;
    std::set<const MSLane*> isActive;
    for (const std::string& id : StringTokenizer(xml.substr(a + 7, b - a - 7)).getVector()) {
        MSLane* lane = MSLane::dictionary(id);
        if (lane != nullptr) {
            active.push_back(lane);
            isActive.insert(lane);
        }
    }
    // MSEdgeControl::patchActiveLanes (MSEdgeControl.cpp:118-133) runs at the begin of the coming step: lanes that got their
    // first vehicle since the last step join in numerical-id order (lanes of multi-lane edges at the front)
    std::map<int, MSLane*> byNumId;
    for (const MSEdge* e : MSEdge::getAllEdges()) {
        for (MSLane* l : e->getLanes()) {
            if (l->getVehicleNumber() > 0 && isActive.count(l) == 0) {
                byNumId[l->getNumericalID()] = l;
            }
        }
    }
    for (const auto& it : byNumId) {
        if (it.second->getEdge().getLanes().size() > 1) {
            active.push_front(it.second);
        } else {
            active.push_back(it.second);
        }
    }
    for (MSLane* lane : active) {
        SumoRNG* rng = lane->getRNG();
        if (clones.count(rng) == 0) {
            clones[rng] = *static_cast<std::mt19937*>(rng);
            p.streams[rng] = std::make_pair(rng->count, 0);
        }
        const MSLane::VehCont& vehs = lane->getVehiclesSecure();
        for (auto it = vehs.rbegin(); it != vehs.rend(); ++it) {
            const double draw = double(clones[rng]() / 4294967296.0);
            p.draws.push_back(std::make_pair((int64_t)(*it)->getNumericalID(), draw));
            p.streams[rng].second++;
        }
        lane->releaseVehicles();
    }
    return p;
}

}  // namespace

#ifndef SUMOB200_NO_MAIN
int
main(int argc, char** argv) {
    std::string flatOut, traceOut, drawsOut, emisOut, tlsSetPhase, vehSetSpeed;
    std::vector<char*> args;
    args.push_back(argv[0]);
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        if (a == "--flat-out" && i + 1 < argc) {
            flatOut = argv[++i];
        } else if (a == "--trace-out" && i + 1 < argc) {
            traceOut = argv[++i];
        } else if (a == "--trajectories-only") {
            gTrajectoriesOnly = true;
        } else if (a == "--draws-out" && i + 1 < argc) {
            drawsOut = argv[++i];
        } else if (a == "--tls-set-phase" && i + 1 < argc) {
            tlsSetPhase = argv[++i];   // "timeMs:tlsID:phase;..." -> libsumo::TrafficLight::setPhase before the step at that time (what a TraCI client does)
        } else if (a == "--veh-set-speed" && i + 1 < argc) {
            vehSetSpeed = argv[++i];   // "timeMs:vehID:speed;..." -> libsumo::Vehicle::setSpeed before the step at that time, if the vehicle is on the road then
        } else if (a == "--emissions-out" && i + 1 < argc) {
            emisOut = argv[++i];   // per step: (t, n) then n x (numerical id, 7 fp64 accumulators of MSDevice_Emissions) as raw bytes
        } else {
            args.push_back(argv[i]);
        }
    }
    static char routeSteps[] = "--route-steps";
    static char zero[] = "0";
    args.push_back(routeSteps);
    args.push_back(zero);
    OptionsCont& oc = OptionsCont::getOptions();
    oc.setApplicationName("sumo", "Eclipse SUMO sumo " VERSION_STRING);
    gSimulation = true;
    int ret = 0;
    MSNet* net = nullptr;
    SUMOTime flowsUntil = -1;
    try {
        XMLSubSys::init();
        OptionsIO::setArgs((int)args.size(), args.data());
        net = NLBuilder::init();
        if (net == nullptr) {
            return 1;
        }
        if (!flatOut.empty()) {
            flowsUntil = expandFlows(net, string2time(oc.getString("begin")), string2time(oc.getString("end")));
            flatten(net, flatOut);
        }
        if (!traceOut.empty()) {
            // lane index as in the flat file
            std::map<const MSLane*, int> laneIdx;
            for (const MSEdge* e : MSEdge::getAllEdges()) {
                for (const MSLane* l : e->getLanes()) {
                    const int idx = (int)laneIdx.size();
                    laneIdx[l] = idx;
                }
            }
            FILE* f = fopen(traceOut.c_str(), "wb");
            FILE* fd = drawsOut.empty() ? nullptr : fopen(drawsOut.c_str(), "wb");
            FILE* fe = emisOut.empty() ? nullptr : fopen(emisOut.c_str(), "wb");
            const SUMOTime begin = string2time(oc.getString("begin"));
            const SUMOTime end = string2time(oc.getString("end"));
            MSNet::SimulationState state = MSNet::SIMSTATE_RUNNING;
            std::multimap<SUMOTime, std::pair<std::string, int> > phaseEvents;
            for (const std::string& ev : StringTokenizer(tlsSetPhase, ";").getVector()) {
                const std::vector<std::string> f = StringTokenizer(ev, ":").getVector();
                if (f.size() == 3) {
                    phaseEvents.insert(std::make_pair((SUMOTime)StringUtils::toLong(f[0]), std::make_pair(f[1], StringUtils::toInt(f[2]))));
                }
            }
            std::multimap<SUMOTime, std::pair<std::string, double> > speedEvents;
            for (const std::string& ev : StringTokenizer(vehSetSpeed, ";").getVector()) {
                const std::vector<std::string> f = StringTokenizer(ev, ":").getVector();
                if (f.size() == 3) {
                    speedEvents.insert(std::make_pair((SUMOTime)StringUtils::toLong(f[0]), std::make_pair(f[1], StringUtils::toDouble(f[2]))));
                }
            }
            // mirror MSNet::simulate (MSNet.cpp:462-500) but record the state after every step
            while (state == MSNet::SIMSTATE_RUNNING) {
                const SUMOTime t = net->getCurrentTimeStep();
                for (auto it = speedEvents.lower_bound(t); it != speedEvents.end() && it->first == t; ++it) {
                    SUMOVehicle* sv = net->getVehicleControl().getVehicle(it->second.first);
                    if (sv != nullptr && sv->isOnRoad()) {
                        libsumo::Vehicle::setSpeed(it->second.first, it->second.second);
                    }
                }
                for (auto it = phaseEvents.lower_bound(t); it != phaseEvents.end() && it->first == t; ++it) {
                    libsumo::TrafficLight::setPhase(it->second.first, it->second.second);
                }
                DrawPrediction pred;
                if (fd != nullptr) {
                    pred = predictDraws(net);
                }
                net->simulationStep();
                if (fd != nullptr) {
                    pred.checkAndWrite(t, fd);
                }
                for (const MSEdge* e : MSEdge::getAllEdges()) {
                    if (e->isVaporizing()) {   // (<vaporizer> starts as an event, nothing at load time tells)
                        throw ProcessError("sumob200: vaporizers are not supported (edge '" + e->getID() + "')");
                    }
                }
                std::vector<const MSVehicle*> on;
                MSVehicleControl& vc = net->getVehicleControl();
                for (auto it = vc.loadedVehBegin(); it != vc.loadedVehEnd(); ++it) {
                    const MSVehicle* v = dynamic_cast<const MSVehicle*>(it->second);
                    if (v != nullptr && v->isOnRoad()) {
                        on.push_back(v);
                    }
                }
                std::sort(on.begin(), on.end(), [](const MSVehicle* a, const MSVehicle* b) {
                    return a->getNumericalID() < b->getNumericalID();
                });
                const int64_t hdr[2] = {(int64_t)t, (int64_t)on.size()};
                fwrite(hdr, 8, 2, f);
                for (const MSVehicle* v : on) {
                    const int64_t ids[2] = {(int64_t)v->getNumericalID(), (int64_t)laneIdx[v->getLane()]};
                    const double st[2] = {v->getPositionOnLane(), v->getSpeed()};
                    fwrite(ids, 8, 2, f);
                    fwrite(st, 8, 2, f);
                }
                if (fe != nullptr) {
                    std::vector<std::pair<int64_t, const MSDevice_Emissions*> > devs;
                    for (const MSVehicle* v : on) {
                        const MSDevice_Emissions* d = static_cast<const MSDevice_Emissions*>(v->getDevice(typeid(MSDevice_Emissions)));
                        if (d != nullptr) {
                            devs.push_back(std::make_pair((int64_t)v->getNumericalID(), d));
                        }
                    }
                    const int64_t ehdr[2] = {(int64_t)t, (int64_t)devs.size()};
                    fwrite(ehdr, 8, 2, fe);
                    for (const auto& d : devs) {
                        const PollutantsInterface::Emissions& em = d.second->myEmissions;
                        const double vals[7] = {em.CO2, em.CO, em.HC, em.fuel, em.NOx, em.PMx, em.electricity};
                        fwrite(&d.first, 8, 1, fe);
                        fwrite(vals, 8, 7, fe);
                    }
                }
                state = net->simulationState(end);
                if (state == MSNet::SIMSTATE_LOADING) {
                    state = MSNet::SIMSTATE_RUNNING;
                }
                if (state == MSNet::SIMSTATE_NO_FURTHER_VEHICLES && net->getCurrentTimeStep() <= flowsUntil) {
                    state = MSNet::SIMSTATE_RUNNING;   // the unexpanded run still has a pending flow here (expandFlows)
                }
            }
            fclose(f);
            if (fd != nullptr) {
                fclose(fd);
            }
            if (fe != nullptr) {
                fclose(fe);
            }
            net->closeSimulation(begin, MSNet::getStateMessage(state));
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
    SystemFrame::close();
    return ret;
}
#endif  // SUMOB200_NO_MAIN
