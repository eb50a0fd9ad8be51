"""ctypes binding of libsumob200.so (include/sumob200.h) -- the host-side mirror of the reference's step-and-query surface.

`Engine.step()` is what `libsumo.simulationStep()` / `MSNet::simulationStep` do for the vehicle update
(src/libsumo/Simulation.cpp:168-188, src/microsim/MSNet.cpp:800-936); `Engine.vehicles()` is the state the
`libsumo.vehicle.get*` getters read (src/libsumo/Vehicle.cpp).  There is no CPU fallback: importing works anywhere, but
creating an engine without the built library or without a CUDA device raises.
"""
import ctypes
import os

import numpy as np

from . import build as _build

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SUMOB200_LIB") or os.path.join(_HERE, "libsumob200.so")   # override: tuning variants only

FLAGS = {
    1: "drive item overflow", 2: "further-lane overflow", 4: "partial occupator overflow", 8: "collision", 16: "teleport",
    32: "emergency stop", 64: "no connection", 128: "unsupported link state", 256: "lane order", 512: "search overflow", 1024: "executeMove pass 2 unresolved", 2048: "look-ahead beyond the partition's ghost column", 4096: "halo message overflow",
    8192: "detector event list / reminder overflow",
}


class SumoB200Error(RuntimeError):
    pass


class _Array(ctypes.Structure):
    _fields_ = [("name", ctypes.c_char_p), ("dtype", ctypes.c_uint32), ("elem_size", ctypes.c_uint32),
                ("count", ctypes.c_uint64), ("data", ctypes.c_void_p)]


class StepResult(ctypes.Structure):
    _fields_ = [("time_ms", ctypes.c_int64), ("running", ctypes.c_int32), ("departed_total", ctypes.c_int32),
                ("arrived_total", ctypes.c_int32), ("collisions", ctypes.c_int32), ("emergency_stops", ctypes.c_int32),
                ("teleports", ctypes.c_int32), ("flags", ctypes.c_int32), ("halting", ctypes.c_int32),
                ("vehicle_updates", ctypes.c_int64), ("teleporting", ctypes.c_int32), ("discarded_total", ctypes.c_int32)]


VEHICLE_STATE = np.dtype([("lane", np.int32), ("route_pos", np.int32), ("pos", np.float64), ("speed", np.float64)])
TRIPINFO = np.dtype([("depart_ms", np.int64), ("depart_delay_ms", np.int64), ("arrival_ms", np.int64), ("duration_ms", np.int64),
                     ("waiting_ms", np.int64), ("time_loss_ms", np.int64), ("depart_lane", np.int32), ("arrival_lane", np.int32),
                     ("waiting_count", np.int32), ("teleport_arrived", np.int32), ("depart_pos", np.float64), ("depart_speed", np.float64),
                     ("arrival_pos", np.float64), ("arrival_speed", np.float64), ("route_length", np.float64), ("speed_factor", np.float64)])
SUMMARY = np.dtype([("time_ms", np.int64), ("inserted", np.int32), ("running", np.int32), ("waiting", np.int32), ("ended", np.int32),
                    ("arrived", np.int32), ("collisions", np.int32), ("teleports", np.int32), ("halting", np.int32), ("stopped", np.int32),
                    ("reserved", np.int32), ("mean_waiting_time", np.float64), ("mean_travel_time", np.float64), ("mean_speed", np.float64),
                    ("mean_speed_relative", np.float64)])
E1_INTERVAL = np.dtype([("n_veh_contrib", np.int32), ("n_veh_entered", np.int32), ("flow", np.float64), ("occupancy", np.float64),
                        ("speed", np.float64), ("harmonic_mean_speed", np.float64), ("length", np.float64)])
E2_INTERVAL = np.dtype([("sampled_seconds", np.float64), ("n_veh_entered", np.int32), ("n_veh_left", np.int32), ("n_veh_seen", np.int32),
                        ("max_jam_length_in_vehicles", np.int32), ("mean_speed", np.float64), ("mean_time_loss", np.float64),
                        ("mean_occupancy", np.float64), ("max_occupancy", np.float64), ("mean_max_jam_length_in_vehicles", np.float64),
                        ("mean_max_jam_length_in_meters", np.float64), ("max_jam_length_in_meters", np.float64),
                        ("jam_length_in_vehicles_sum", np.int32), ("started_halts", np.int32), ("jam_length_in_meters_sum", np.float64),
                        ("mean_halting_duration", np.float64), ("max_halting_duration", np.float64), ("halting_duration_sum", np.float64),
                        ("mean_interval_halting_duration", np.float64), ("max_interval_halting_duration", np.float64),
                        ("interval_halting_duration_sum", np.float64), ("mean_vehicle_number", np.float64),
                        ("max_vehicle_number", np.int32), ("reserved", np.int32)])
LANE_MEANDATA = np.dtype([("sample_seconds", np.float64), ("travelled_distance", np.float64), ("wait_seconds", np.float64), ("time_loss", np.float64),
                          ("front_sample_seconds", np.float64), ("front_travelled_distance", np.float64), ("veh_length_sum", np.float64),
                          ("occupation_sum", np.float64), ("minimal_vehicle_length", np.float64), ("departed", np.int32), ("arrived", np.int32),
                          ("entered", np.int32), ("left", np.int32), ("teleported", np.int32), ("reserved", np.int32)])
LANE_STATS = np.dtype([("vehicle_number", np.int32), ("halting_number", np.int32), ("mean_speed", np.float64), ("occupancy", np.float64),
                       ("waiting_time", np.float64)])
FCD_ROW = np.dtype([("slot", np.int32), ("lane", np.int32), ("pos", np.float64), ("speed", np.float64),
                    ("x", np.float64), ("y", np.float64), ("angle", np.float64)])

_DTYPE_CODES = {np.dtype(np.uint8): 0, np.dtype(np.int32): 1, np.dtype(np.int64): 2, np.dtype(np.float64): 3, np.dtype(np.uint64): 4}

_lib = None


def load_library(path=None):
    """dlopen libsumob200.so and declare the prototypes of include/sumob200.h"""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise SumoB200Error("%s is missing: run `python -m sumo_b200.build` (needs nvcc); there is no CPU fallback" % path)
    lib = ctypes.CDLL(path)
    P = ctypes.POINTER
    lib.sumob200_create.argtypes = [P(_Array), ctypes.c_int, ctypes.c_int, P(ctypes.c_void_p)]
    lib.sumob200_create_from_file.argtypes = [ctypes.c_char_p, ctypes.c_int, P(ctypes.c_void_p)]
    lib.sumob200_destroy.argtypes = [ctypes.c_void_p]
    lib.sumob200_destroy.restype = None
    lib.sumob200_step.argtypes = [ctypes.c_void_p, ctypes.c_int, P(StepResult)]
    lib.sumob200_step_timed.argtypes = [ctypes.c_void_p, ctypes.c_int, P(StepResult), P(ctypes.c_double)]
    lib.sumob200_mirror_all.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    lib.sumob200_fcd_snapshot.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, P(ctypes.c_int)]
    lib.sumob200_fcd_format_xml.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_char_p, ctypes.c_uint64, P(ctypes.c_uint64)]
    lib.sumob200_fcd_begin.argtypes = [ctypes.c_void_p]
    lib.sumob200_fcd_end.argtypes = [ctypes.c_void_p, P(ctypes.c_void_p), P(ctypes.c_int)]
    lib.sumob200_tripinfo_all.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    lib.sumob200_summary_now.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.sumob200_summary_now_static.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.sumob200_meandata_interval.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    lib.sumob200_lane_stats_all.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    lib.sumob200_e1_count.argtypes = [ctypes.c_void_p]
    lib.sumob200_e2_count.argtypes = [ctypes.c_void_p]
    lib.sumob200_e2_interval.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    lib.sumob200_e1_interval.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int]
    lib.sumob200_emissions.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64]
    lib.sumob200_num_tls.argtypes = [ctypes.c_void_p]
    lib.sumob200_vehicles_set_speed.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    lib.sumob200_tls_set_phase.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
    lib.sumob200_tls_get_phase.argtypes = [ctypes.c_void_p, ctypes.c_int, P(ctypes.c_int), P(ctypes.c_int64)]
    lib.sumob200_num_vehicles.argtypes = [ctypes.c_void_p]
    lib.sumob200_num_lanes.argtypes = [ctypes.c_void_p]
    lib.sumob200_time_ms.argtypes = [ctypes.c_void_p]
    lib.sumob200_time_ms.restype = ctypes.c_int64
    lib.sumob200_kernel_launches.argtypes = [ctypes.c_void_p]
    lib.sumob200_kernel_launches.restype = ctypes.c_int64
    lib.sumob200_set_random_draws.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    lib.sumob200_set_random_mode.argtypes = [ctypes.c_void_p, ctypes.c_int]
    lib.sumob200_set_profiling.argtypes = [ctypes.c_void_p, ctypes.c_int]
    lib.sumob200_set_reslot_period.argtypes = [ctypes.c_void_p, ctypes.c_int]
    lib.sumob200_phase_times.argtypes = [ctypes.c_void_p, P(ctypes.c_double), ctypes.c_int, ctypes.c_int]
    lib.sumob200_last_error.restype = ctypes.c_char_p
    lib.sumob200_version.restype = ctypes.c_char_p
    if path == LIB_PATH:
        _lib = lib
    return lib


def _check(lib, rc, what):
    if rc != 0:
        raise SumoB200Error("%s failed (%d): %s" % (what, rc, lib.sumob200_last_error().decode()))


class Engine:
    """one simulation on one GPU"""

    def __init__(self, scenario, device=0):
        """scenario: path of a .flat file, or {name: np.ndarray} as returned by flatfile.load_flat / tiler.tile"""
        self._lib = load_library()
        self._h = ctypes.c_void_p()
        if isinstance(scenario, (str, bytes, os.PathLike)):
            rc = self._lib.sumob200_create_from_file(os.fsencode(scenario), device, ctypes.byref(self._h))
        else:
            keep = []
            arr = (_Array * len(scenario))()
            for i, (name, a) in enumerate(scenario.items()):
                a = np.ascontiguousarray(a)
                keep.append(a)
                arr[i].name = name.encode()
                arr[i].dtype = _DTYPE_CODES[a.dtype]
                arr[i].elem_size = a.dtype.itemsize
                arr[i].count = a.size
                arr[i].data = a.ctypes.data
            rc = self._lib.sumob200_create(arr, len(scenario), device, ctypes.byref(self._h))
        _check(self._lib, rc, "sumob200_create")
        self.num_vehicles = self._lib.sumob200_num_vehicles(self._h)
        self.num_lanes = self._lib.sumob200_num_lanes(self._h)
        self._state = np.zeros(max(self.num_vehicles, 1), dtype=VEHICLE_STATE)

    def close(self):
        if self._h:
            self._lib.sumob200_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    @property
    def time_ms(self):
        return self._lib.sumob200_time_ms(self._h)

    @property
    def kernel_launches(self):
        return self._lib.sumob200_kernel_launches(self._h)

    def step(self, n=1, result=True):
        """run n simulation steps; returns the StepResult after the last one (None if result is False: no host sync)"""
        if result:
            r = StepResult()
            _check(self._lib, self._lib.sumob200_step(self._h, n, ctypes.byref(r)), "sumob200_step")
            return r
        _check(self._lib, self._lib.sumob200_step(self._h, n, None), "sumob200_step")
        return None

    def step_timed(self, n):
        """run n steps; returns (StepResult, device milliseconds measured with CUDA events on the engine's stream)"""
        r = StepResult()
        ms = ctypes.c_double()
        _check(self._lib, self._lib.sumob200_step_timed(self._h, n, ctypes.byref(r), ctypes.byref(ms)), "sumob200_step_timed")
        return r, ms.value

    def vehicles(self, out=None):
        """state of all vehicle slots as a structured array (lane, route_pos, pos, speed); lane < 0 = not on the road"""
        buf = self._state if out is None else out
        _check(self._lib, self._lib.sumob200_mirror_all(self._h, buf.ctypes.data, buf.shape[0]), "sumob200_mirror_all")
        return buf[:self.num_vehicles]

    def fcd(self):
        """FCD rows (slot, lane, pos, speed, x, y, angle[navi degrees]) of the vehicles on the road, ordered by slot"""
        rows = np.zeros(max(self.num_vehicles, 1), dtype=FCD_ROW)
        n = ctypes.c_int()
        _check(self._lib, self._lib.sumob200_fcd_snapshot(self._h, rows.ctypes.data, rows.shape[0], ctypes.byref(n)), "sumob200_fcd_snapshot")
        return rows[:n.value]

    def fcd_xml(self, rows, time_ms):
        """the <timestep> element of the reference's --fcd-output for these rows (bytes)"""
        rows = np.ascontiguousarray(rows, dtype=FCD_ROW)
        n = ctypes.c_uint64()
        cap = 160 * len(rows) + 256
        buf = ctypes.create_string_buffer(cap)
        _check(self._lib, self._lib.sumob200_fcd_format_xml(self._h, rows.ctypes.data, len(rows), time_ms, buf, cap, ctypes.byref(n)), "sumob200_fcd_format_xml")
        if n.value > cap:
            cap = n.value
            buf = ctypes.create_string_buffer(cap)
            _check(self._lib, self._lib.sumob200_fcd_format_xml(self._h, rows.ctypes.data, len(rows), time_ms, buf, cap, ctypes.byref(n)), "sumob200_fcd_format_xml")
        return buf.raw[:n.value]

    def fcd_begin(self):
        """start the device-side compaction + pinned async D2H of the FCD rows; the next step may be issued right away"""
        _check(self._lib, self._lib.sumob200_fcd_begin(self._h), "sumob200_fcd_begin")

    def fcd_end(self):
        """wait for the copy started by fcd_begin; returns a view of the engine's pinned buffer (valid until the second next fcd_begin)"""
        p = ctypes.c_void_p()
        n = ctypes.c_int()
        _check(self._lib, self._lib.sumob200_fcd_end(self._h, ctypes.byref(p), ctypes.byref(n)), "sumob200_fcd_end")
        if n.value == 0:
            return np.zeros(0, dtype=FCD_ROW)
        buf = (ctypes.c_char * (n.value * FCD_ROW.itemsize)).from_address(p.value)
        return np.frombuffer(buf, dtype=FCD_ROW, count=n.value)

    def tripinfo(self):
        """trip record of every vehicle slot (MSDevice_Tripinfo::generateOutput fields), structured array TRIPINFO"""
        out = np.zeros(max(self.num_vehicles, 1), dtype=TRIPINFO)
        _check(self._lib, self._lib.sumob200_tripinfo_all(self._h, out.ctypes.data, out.shape[0]), "sumob200_tripinfo_all")
        return out[:self.num_vehicles]

    def summary(self):
        """the attributes of the <step> line MSNet::writeSummaryOutput writes for the step executed last (record of dtype SUMMARY)"""
        out = np.zeros(1, dtype=SUMMARY)
        _check(self._lib, self._lib.sumob200_summary_now(self._h, out.ctypes.data), "sumob200_summary_now")
        return out[0]

    def summary_static(self):
        """summary() through sumob200_summary_now_static: the same record, computed without any device allocation (scratch in a
        static device array of the library; at most SUMOB200_SUMMARY_STATIC_MAX_VEHICLES = 2,097,152 vehicle slots)"""
        out = np.zeros(1, dtype=SUMMARY)
        _check(self._lib, self._lib.sumob200_summary_now_static(self._h, out.ctypes.data), "sumob200_summary_now_static")
        return out[0]

    def lane_stats(self):
        """libsumo.lane.getLastStepVehicleNumber / HaltingNumber / MeanSpeed / Occupancy / getWaitingTime of every lane (dtype LANE_STATS)"""
        out = np.zeros(max(self.num_lanes, 1), dtype=LANE_STATS)
        _check(self._lib, self._lib.sumob200_lane_stats_all(self._h, out.ctypes.data, out.shape[0]), "sumob200_lane_stats_all")
        return out[:self.num_lanes]

    def meandata_interval(self):
        """the raw MSLaneMeanDataValues members of every lane since the last call (dtype LANE_MEANDATA, indexed by lane); resets them"""
        out = np.zeros(max(self.num_lanes, 1), dtype=LANE_MEANDATA)
        _check(self._lib, self._lib.sumob200_meandata_interval(self._h, out.ctypes.data, out.shape[0]), "sumob200_meandata_interval")
        return out[:self.num_lanes]

    def e1_interval(self, begin_ms, end_ms):
        """MSInductLoop::writeXMLOutput for every E1 detector of the scenario: the aggregates of [begin_ms, end_ms) -- end_ms must be
        the engine's current time --, then the detectors are reset (array of dtype E1_INTERVAL, one record per detector)"""
        n = self._lib.sumob200_e1_count(self._h)
        out = np.zeros(max(n, 1), dtype=E1_INTERVAL)
        _check(self._lib, self._lib.sumob200_e1_interval(self._h, int(begin_ms), int(end_ms), out.ctypes.data, out.shape[0]), "sumob200_e1_interval")
        return out[:n]

    def e2_interval(self):
        """MSE2Collector::writeXMLOutput for every E2 detector of the scenario (the interval that ends now), then the detectors are
        reset (array of dtype E2_INTERVAL, one record per detector)"""
        n = self._lib.sumob200_e2_count(self._h)
        out = np.zeros(max(n, 1), dtype=E2_INTERVAL)
        _check(self._lib, self._lib.sumob200_e2_interval(self._h, out.ctypes.data, out.shape[0]), "sumob200_e2_interval")
        return out[:n]

    def emissions(self):
        """[7, num_vehicles] accumulated CO2, CO, HC, fuel, NOx, PMx, electricity of every vehicle slot"""
        out = np.zeros((7, self.num_vehicles))
        _check(self._lib, self._lib.sumob200_emissions(self._h, out.ctypes.data, out.size), "sumob200_emissions")
        return out

    def set_speed(self, slots, speeds):
        """libsumo.vehicle.setSpeed for a batch of vehicle slots (speed < 0 = back to the car-following model)"""
        s = np.ascontiguousarray(np.atleast_1d(slots), dtype=np.int32)
        v = np.ascontiguousarray(np.atleast_1d(speeds), dtype=np.float64)
        assert s.shape == v.shape
        _check(self._lib, self._lib.sumob200_vehicles_set_speed(self._h, s.ctypes.data, v.ctypes.data, s.size), "sumob200_vehicles_set_speed")

    @property
    def num_tls(self):
        return self._lib.sumob200_num_tls(self._h)

    def tls_set_phase(self, tls, phase):
        """libsumo.trafficlight.setPhase between two steps; tls = index in the scenario's tl_names"""
        _check(self._lib, self._lib.sumob200_tls_set_phase(self._h, int(tls), int(phase)), "sumob200_tls_set_phase")

    def tls_phase(self, tls):
        """(current phase index, next switch time in ms) = trafficlight.getPhase / getNextSwitch"""
        ph, ns = ctypes.c_int(), ctypes.c_int64()
        _check(self._lib, self._lib.sumob200_tls_get_phase(self._h, int(tls), ctypes.byref(ph), ctypes.byref(ns)), "sumob200_tls_get_phase")
        return ph.value, ns.value

    def set_random_draws(self, draws):
        d = np.ascontiguousarray(draws, dtype=np.float64)
        _check(self._lib, self._lib.sumob200_set_random_draws(self._h, d.ctypes.data, d.size), "sumob200_set_random_draws")

    def set_random_mode(self, mode):
        _check(self._lib, self._lib.sumob200_set_random_mode(self._h, mode), "sumob200_set_random_mode")

    def set_reslot_period(self, steps):
        """renumber the on-road vehicles in lane order every `steps` steps (memory locality; results unchanged; 0 = never)"""
        _check(self._lib, self._lib.sumob200_set_reslot_period(self._h, int(steps)), "sumob200_set_reslot_period")

    def set_profiling(self, enabled):
        _check(self._lib, self._lib.sumob200_set_profiling(self._h, 1 if enabled else 0), "sumob200_set_profiling")

    def phase_times(self, reset=True):
        ms = (ctypes.c_double * 5)()
        _check(self._lib, self._lib.sumob200_phase_times(self._h, ms, 5, 1 if reset else 0), "sumob200_phase_times")
        return dict(zip(("pre", "plan", "execute", "handoff", "insert"), list(ms)))


def describe_flags(flags):
    return [name for bit, name in FLAGS.items() if flags & bit]


def ensure_built():
    return _build.build()
