"""ctypes binding of the lsb_* C ABI (include/lidarslam_b200.h).

This is the Python-side mirror used by tests and bench.py; the product itself is
the C-ABI shared library (CUDA kernels + C++ host driver).  `CApi(lib, prefix)`
is generic over the symbol prefix so that the CPU oracle (oracle/pyorc.py, test
infrastructure) can be driven through the very same calls with prefix "orc_".
Nothing in this package imports the oracle.
"""
import ctypes as C
import os

import numpy as np

# bit-identical to LidarSlam::LidarPoint (reference LidarPoint.h:31-64)
POINT_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("w", "<f4"), ("time", "<f8"), ("intensity", "<f4"),
     ("laser_id", "<u2"), ("device_id", "u1"), ("label", "u1")], align=True)
assert POINT_DTYPE.itemsize == 32

# raw driver clouds of the input adaptors (reference ros_wrapping/lidar_conversions)
VELODYNE_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("_pad0", "<f4"), ("intensity", "<f4"), ("ring", "<u2"),
                           ("_pad1", "<u2"), ("time", "<f4"), ("_pad2", "<f4")])      # velodyne_pcl::PointXYZIRT
XYZI_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("_pad0", "<f4"), ("intensity", "<f4"), ("_pad1", "<f4", (3,))])  # pcl::PointXYZI
assert VELODYNE_DTYPE.itemsize == 32 and XYZI_DTYPE.itemsize == 32

EDGE, PLANE, BLOB = 0, 1, 2
PCD_ASCII, PCD_BINARY, PCD_BINARY_COMPRESSED = 0, 1, 2   # reference PointCloudStorage.h:60-65
KEYPOINT_TYPES = (EDGE, PLANE, BLOB)
MATCH_STATUS_NAMES = ("SUCCESS", "BAD_MODEL_PARAMETRIZATION", "NOT_ENOUGH_NEIGHBORS", "NEIGHBORS_TOO_FAR",
                      "BAD_PCA_STRUCTURE", "INVALID_NUMERICAL", "MSE_TOO_LARGE", "UNKNOWN")


class KeParams(C.Structure):
    _fields_ = [("neighbor_width", C.c_int32), ("min_distance_to_sensor", C.c_float), ("min_beam_surface_angle", C.c_float),
                ("plane_sin_angle_threshold", C.c_float), ("edge_sin_angle_threshold", C.c_float),
                ("dist_to_line_threshold", C.c_float), ("edge_depth_gap_threshold", C.c_float),
                ("edge_saliency_threshold", C.c_float), ("edge_intensity_gap_threshold", C.c_float),
                ("azimuthal_resolution", C.c_float)]


class MatchParams(C.Structure):
    _fields_ = [("single_edge_per_ring", C.c_int32), ("max_neighbors_distance", C.c_double),
                ("edge_nb_neighbors", C.c_uint32), ("edge_min_nb_neighbors", C.c_uint32), ("edge_max_model_error", C.c_double),
                ("plane_nb_neighbors", C.c_uint32), ("planarity_threshold", C.c_double), ("plane_max_model_error", C.c_double),
                ("blob_nb_neighbors", C.c_uint32), ("saturation_distance", C.c_double)]


class Params(C.Structure):
    _fields_ = [("nb_threads", C.c_int32), ("use_keypoints", C.c_int32 * 3), ("ego_motion", C.c_int32),
                ("undistortion", C.c_int32), ("verbosity", C.c_int32), ("logging_timeout", C.c_double),
                ("two_d_mode", C.c_int32),
                ("ego_motion_icp_max_iter", C.c_uint32), ("localization_icp_max_iter", C.c_uint32),
                ("ego_motion_lm_max_iter", C.c_uint32), ("localization_lm_max_iter", C.c_uint32),
                ("ego_motion_max_neighbors_distance", C.c_double), ("localization_max_neighbors_distance", C.c_double),
                ("ego_motion_edge_nb_neighbors", C.c_uint32), ("ego_motion_edge_min_nb_neighbors", C.c_uint32),
                ("ego_motion_edge_max_model_error", C.c_double),
                ("localization_edge_nb_neighbors", C.c_uint32), ("localization_edge_min_nb_neighbors", C.c_uint32),
                ("localization_edge_max_model_error", C.c_double),
                ("ego_motion_plane_nb_neighbors", C.c_uint32), ("ego_motion_planarity_threshold", C.c_double),
                ("ego_motion_plane_max_model_error", C.c_double),
                ("localization_plane_nb_neighbors", C.c_uint32), ("localization_planarity_threshold", C.c_double),
                ("localization_plane_max_model_error", C.c_double),
                ("localization_blob_nb_neighbors", C.c_uint32),
                ("ego_motion_init_saturation_distance", C.c_double), ("ego_motion_final_saturation_distance", C.c_double),
                ("localization_init_saturation_distance", C.c_double), ("localization_final_saturation_distance", C.c_double),
                ("kf_distance_threshold", C.c_double), ("kf_angle_threshold", C.c_double),
                ("map_update", C.c_int32), ("max_extrapolation_ratio", C.c_double),
                ("min_nb_matched_keypoints", C.c_uint32), ("overlap_sampling_ratio", C.c_float),
                ("time_window_duration", C.c_float), ("velocity_limits", C.c_float * 2), ("acceleration_limits", C.c_float * 2)]


class ConversionParams(C.Structure):
    _fields_ = [("laser_id_mapping", C.c_void_p), ("n_laser_id_mapping", C.c_int32), ("device_id", C.c_int32), ("rpm", C.c_double),
                ("timestamp_first_packet", C.c_int32)]


class RegistrationError(C.Structure):
    _fields_ = [("covariance", C.c_double * 36), ("position_error", C.c_double),
                ("position_error_direction", C.c_double * 3), ("orientation_error", C.c_double),
                ("orientation_error_direction", C.c_double * 3)]


class FrameStats(C.Structure):
    _fields_ = [("n_points", C.c_uint32), ("n_keypoints", C.c_uint32 * 3), ("n_submap", C.c_uint32 * 3),
                ("ego_icp_iters", C.c_uint32), ("loc_icp_iters", C.c_uint32),
                ("ego_lm_iters", C.c_uint32), ("loc_lm_iters", C.c_uint32),
                ("n_matches", C.c_uint32 * 3), ("ego_n_matches", C.c_uint32 * 3),
                ("knn_queries", C.c_uint32), ("is_keyframe", C.c_uint32), ("n_map", C.c_uint32 * 3),
                ("kernel_launches", C.c_uint32),
                ("ms_extract", C.c_float), ("ms_ego", C.c_float), ("ms_localization", C.c_float),
                ("ms_map_update", C.c_float), ("ms_total", C.c_float), ("ms_knn", C.c_float),
                ("knn_launches", C.c_uint32), ("ms_fit", C.c_float), ("ms_lm", C.c_float), ("lm_launches", C.c_uint32),
                ("d2h_bytes", C.c_uint64), ("h2d_bytes", C.c_uint64),
                ("map_update_deferred", C.c_uint32), ("prefetched", C.c_uint32), ("launches_total", C.c_uint32)]


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def as_points(a):
    a = np.ascontiguousarray(a)
    assert a.dtype == POINT_DTYPE, a.dtype
    return a


def _mat16(T):
    return np.ascontiguousarray(np.asarray(T, dtype=np.float64).reshape(16))


class CApi:
    """One SLAM handle behind a C ABI with the given symbol prefix.

    Method names follow LidarSlam::Slam (reference Slam.h:98-394)."""

    def __init__(self, lib, prefix, create_args=()):
        self.lib, self.px = lib, prefix
        f = self._f("create")
        if prefix == "lsb_":
            h = C.c_void_p()
            f.restype = C.c_int
            rc = f(C.byref(h), C.c_int(create_args[0] if create_args else 0))
            if rc != 0:
                raise RuntimeError("lsb_create failed (%d): %s" % (rc, self.last_error()))
            self.h = h
        else:
            f.restype = C.c_void_p
            self.h = C.c_void_p(f())
        self._keep = None

    def _f(self, name):
        return getattr(self.lib, self.px + name)

    def last_error(self):
        if self.px != "lsb_":
            return ""
        f = self.lib.lsb_last_error
        f.restype = C.c_char_p
        return (f() or b"").decode()

    def _check(self, rc):
        if rc < 0:
            raise RuntimeError("%s call failed (%d): %s" % (self.px, rc, self.last_error()))
        return rc

    def close(self):
        if self.h:
            self._f("destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- parameters --------------------------------------------------------
    def GetParams(self):
        p = Params()
        self._check(self._f("get_params")(self.h, C.byref(p)))
        return p

    def SetParams(self, p):
        self._check(self._f("set_params")(self.h, C.byref(p)))

    def GetExtractorParams(self, device_id=0):
        p = KeParams()
        self._check(self._f("get_extractor_params")(self.h, C.c_int(device_id), C.byref(p)))
        return p

    def SetExtractorParams(self, p, device_id=0):
        self._check(self._f("set_extractor_params")(self.h, C.c_int(device_id), C.byref(p)))

    def SetBaseToLidarOffset(self, T, device_id=0):
        self._check(self._f("set_base_to_lidar")(self.h, C.c_int(device_id), _ptr(_mat16(T))))

    def SetVoxelGridLeafSize(self, k, size):
        self._check(self._f("set_voxel_grid_leaf_size")(self.h, C.c_int(k), C.c_double(size)))

    def SetVoxelGridSize(self, size):
        self._check(self._f("set_voxel_grid_size")(self.h, C.c_int(size)))

    def SetVoxelGridResolution(self, res):
        self._check(self._f("set_voxel_grid_resolution")(self.h, C.c_double(res)))

    def SetVoxelGridDecayingThreshold(self, seconds):
        self._check(self._f("set_voxel_grid_decaying_threshold")(self.h, C.c_double(seconds)))

    def SetVoxelGridMinFramesPerVoxel(self, n):
        self._check(self._f("set_voxel_grid_min_frames_per_voxel")(self.h, C.c_uint(n)))

    def SetVoxelGridSamplingMode(self, k, mode):
        self._check(self._f("set_voxel_grid_sampling_mode")(self.h, C.c_int(k), C.c_int(mode)))

    def Reset(self, reset_log=True):
        self._check(self._f("reset")(self.h, C.c_int(1 if reset_log else 0)))

    def ClearMaps(self):
        self._check(self._f("clear_maps")(self.h))

    def SetWorldTransformFromGuess(self, T):
        self._check(self._f("set_world_transform_from_guess")(self.h, _ptr(_mat16(T))))

    # ---- per-frame entry ---------------------------------------------------
    def AddFrames(self, frames, stamps_us, seqs=None):
        frames = [as_points(f) for f in frames]
        n = len(frames)
        ptrs = (C.c_void_p * n)(*[f.ctypes.data for f in frames])
        sizes = (C.c_size_t * n)(*[len(f) for f in frames])
        stamps = (C.c_uint64 * n)(*[int(s) for s in stamps_us])
        sq = (C.c_uint32 * n)(*[int(s) for s in (seqs or [0] * n)])
        self._keep = frames  # the reference retains the input clouds until the next call (Slam.cxx:237)
        f = self._f("add_frames")
        f.restype = C.c_int
        return self._check(f(self.h, C.c_int(n), ptrs, sizes, stamps, sq))

    def AddFrame(self, frame, stamp_us, seq=0):
        return self.AddFrames([frame], [stamp_us], [seq])

    # ---- outputs -----------------------------------------------------------
    def GetWorldTransform(self):
        T = np.zeros(16)
        t = C.c_double()
        self._check(self._f("get_world_transform")(self.h, _ptr(T), C.byref(t)))
        return T.reshape(4, 4), t.value

    def GetTrajectory(self):
        """Slam::GetTrajectory: (count, poses [n, 4, 4], times [n]) of the logged states."""
        f = self._f("get_trajectory")
        f.restype = C.c_long
        n = int(f(self.h, None, None, C.c_long(0)))
        T, t = np.zeros((n, 16)), np.zeros(n)
        if n:
            f(self.h, _ptr(T), _ptr(t), C.c_long(n))
        return n, T.reshape(n, 4, 4), t

    def GetRelativeTransform(self):
        T = np.zeros(16)
        self._check(self._f("get_relative_transform")(self.h, _ptr(T)))
        return T.reshape(4, 4)

    def GetTransformCovariance(self):
        c = np.zeros(36)
        self._check(self._f("get_covariance")(self.h, _ptr(c)))
        return c

    def GetRegistrationError(self):
        e = RegistrationError()
        self._check(self._f("get_registration_error")(self.h, C.byref(e)))
        return e

    def _cloud(self, name, *args):
        f = self._f(name)
        f.restype = C.c_long
        n = f(self.h, *args, None, C.c_long(0))
        if n < 0:
            self._check(int(n))
        out = np.zeros(int(n), dtype=POINT_DTYPE)
        if n:
            f(self.h, *args, _ptr(out), C.c_long(n))
        return out

    def GetKeypoints(self, k, world=False):
        return self._cloud("get_keypoints", C.c_int(k), C.c_int(1 if world else 0))

    def GetMap(self, k, clean=False):
        return self._cloud("get_map", C.c_int(k), C.c_int(1 if clean else 0))

    # ---- input adaptors (reference ros_wrapping/lidar_conversions) ----------
    def _conversion(self, name, pts, extra, laser_id_mapping, device_id, rpm, timestamp_first_packet):
        m = np.ascontiguousarray(laser_id_mapping, dtype=np.int32) if laser_id_mapping is not None else None
        cp = ConversionParams(_ptr(m) if m is not None else None, len(m) if m is not None else 0, device_id, rpm, 1 if timestamp_first_packet else 0)
        f = self._f(name)
        f.restype = C.c_long
        n = f(self.h, _ptr(pts), *extra, C.byref(cp), None, C.c_long(0))
        if n < 0:
            self._check(int(n))
        out = np.zeros(int(n), dtype=POINT_DTYPE)
        if n:
            f(self.h, _ptr(pts), *extra, C.byref(cp), _ptr(out), C.c_long(n))
        return out

    def ConvertVelodyne(self, pts, laser_id_mapping=None, device_id=0, rpm=600.0, timestamp_first_packet=False):
        """VelodyneToLidarNode::Callback; the converted frame also stays inside the handle for AddConvertedFrame."""
        pts = np.ascontiguousarray(pts)
        assert pts.dtype == VELODYNE_DTYPE
        return self._conversion("convert_velodyne", pts, (C.c_size_t(len(pts)),), laser_id_mapping, device_id, rpm, timestamp_first_packet)

    def ConvertRobosense(self, pts, width, height, laser_id_mapping=None, device_id=0, rpm=600.0):
        """RobosenseToLidarNode::Callback on an organised width x height cloud (ring-major)."""
        pts = np.ascontiguousarray(pts)
        assert pts.dtype == XYZI_DTYPE and len(pts) == width * height
        return self._conversion("convert_robosense", pts, (C.c_uint32(width), C.c_uint32(height)), laser_id_mapping, device_id, rpm, False)

    def AddConvertedFrame(self, stamp_us, seq=0):
        return self._check(self._f("add_converted_frame")(self.h, C.c_uint64(stamp_us), C.c_uint32(seq)))

    # ---- external sensor constraints (reference Slam.h:331-344) ------------
    def AddWheelOdomMeasurement(self, time, distance):
        m = (C.c_double * 2)(time, distance)
        self._check(self._f("add_wheel_odom_measurement")(self.h, m))

    def AddGravityMeasurement(self, time, acceleration):
        m = (C.c_double * 4)(time, *[float(v) for v in acceleration])
        self._check(self._f("add_gravity_measurement")(self.h, m))

    def ClearSensorMeasurements(self):
        self._check(self._f("clear_sensor_measurements")(self.h))

    def SetWheelOdomWeight(self, w):
        self._check(self._f("set_wheel_odom_weight")(self.h, C.c_double(w)))

    def SetGravityWeight(self, w):
        self._check(self._f("set_gravity_weight")(self.h, C.c_double(w)))

    def SetSensorTimeOffset(self, t):
        self._check(self._f("set_sensor_time_offset")(self.h, C.c_double(t)))

    def GetGravityReference(self):
        g = np.zeros(3)
        self._check(self._f("get_gravity_reference")(self.h, _ptr(g)))
        return g

    def SaveMapsToPCD(self, file_prefix, pcd_format=PCD_BINARY_COMPRESSED, filtered=True):
        """Slam::SaveMapsToPCD (reference Slam.cxx:504-517)."""
        f = self._f("save_maps_to_pcd")
        f.restype = C.c_int
        return self._check(f(self.h, os.fsencode(file_prefix), C.c_int(pcd_format), C.c_int(1 if filtered else 0)))

    def LoadMapsFromPCD(self, file_prefix, reset_maps=True):
        """Slam::LoadMapsFromPCD (reference Slam.cxx:520-543); returns the bit mask of the keypoint types loaded."""
        f = self._f("load_maps_from_pcd")
        f.restype = C.c_int
        return self._check(f(self.h, os.fsencode(file_prefix), C.c_int(1 if reset_maps else 0)))

    def GetTargetSubMap(self, k):
        return self._cloud("get_submap", C.c_int(k))

    def GetRegisteredFrame(self):
        return self._cloud("get_registered_frame")

    def GetMatchDebug(self, stage, k):
        f = self._f("get_match_debug")
        f.restype = C.c_long
        n = int(f(self.h, C.c_int(stage), C.c_int(k), None, None, C.c_long(0)))
        st = np.zeros(n, dtype=np.uint8)
        w = np.zeros(n, dtype=np.float64)
        if n:
            f(self.h, C.c_int(stage), C.c_int(k), _ptr(st), _ptr(w), C.c_long(n))
        return st, w

    def GetKeDebug(self, field, device_id=0):
        f = self._f("get_ke_debug")
        f.restype = C.c_long
        n = int(f(self.h, C.c_int(device_id), C.c_int(field), None, C.c_long(0)))
        out = np.zeros(max(n, 0), dtype=np.float32)
        if n > 0:
            f(self.h, C.c_int(device_id), C.c_int(field), _ptr(out), C.c_long(n))
        return out

    def GetFrameStats(self):
        s = FrameStats()
        self._check(self._f("get_frame_stats")(self.h, C.byref(s)))
        return s

    def GetOverlapEstimation(self):
        f = self._f("get_overlap")
        f.restype = C.c_float
        return float(f(self.h))

    def GetComplyMotionLimits(self):
        return bool(self._f("get_comply_motion_limits")(self.h))

    def GetTotalMatchedKeypoints(self):
        f = self._f("get_total_matched_keypoints")
        f.restype = C.c_uint
        return int(f(self.h))

    # ---- kernel-level entry points ------------------------------------------
    def ExtractKeypoints(self, pts):
        pts = as_points(pts)
        labels = np.zeros(len(pts), dtype=np.uint8)
        counts = (C.c_uint32 * 3)()
        self._check(self._f("extract_keypoints")(self.h, _ptr(pts), C.c_size_t(len(pts)), _ptr(labels), counts))
        kps = [self._cloud("get_extracted_keypoints", C.c_int(k)) for k in KEYPOINT_TYPES]
        return labels, list(counts), kps

    def KnnSearch(self, cloud, queries, k):
        cloud = as_points(cloud)
        q = np.ascontiguousarray(queries, dtype=np.float32).reshape(-1, 3)
        idx = np.zeros((len(q), k), dtype=np.int32)
        d2 = np.zeros((len(q), k), dtype=np.float32)
        self._check(self._f("knn_search")(self.h, _ptr(cloud), C.c_size_t(len(cloud)), _ptr(q), C.c_size_t(len(q)),
                                          C.c_int(k), _ptr(idx), _ptr(d2)))
        return idx, d2

    def MatchAndSolve(self, keypoints, targets, mp, lm_max_iter, prior):
        kps = [as_points(a) for a in keypoints]
        tgs = [as_points(a) for a in targets]
        kp_ptrs = (C.c_void_p * 3)(*[a.ctypes.data for a in kps])
        tg_ptrs = (C.c_void_p * 3)(*[a.ctypes.data for a in tgs])
        nk = (C.c_size_t * 3)(*[len(a) for a in kps])
        nt = (C.c_size_t * 3)(*[len(a) for a in tgs])
        total = sum(len(a) for a in kps)
        status = np.zeros(total, dtype=np.uint8)
        weight = np.zeros(total, dtype=np.float64)
        pose = np.zeros(16)
        acc, its, cost = C.c_int(), C.c_int(), C.c_double()
        err = RegistrationError()
        rc = self._check(self._f("match_and_solve")(self.h, kp_ptrs, nk, tg_ptrs, nt, C.byref(mp), C.c_int(lm_max_iter),
                                                    _ptr(_mat16(prior)), _ptr(status), _ptr(weight), _ptr(pose),
                                                    C.byref(acc), C.byref(its), C.byref(cost), C.byref(err)))
        return dict(rc=rc, status=status, weight=weight, pose=pose.reshape(4, 4), accepted=acc.value,
                    iterations=its.value, cost=cost.value, err=err)

    def MapLoadPoints(self, k, pts, fixed=False):
        """Body of Slam::LoadMapsFromPCD for one keypoint type (reference Slam.cxx:536-537) on host points."""
        pts = as_points(pts)
        self._check(self._f("map_load_points")(self.h, C.c_int(k), _ptr(pts), C.c_long(len(pts)), C.c_int(1 if fixed else 0)))

    def MapAdd(self, k, pts, fixed=False, current_time=0.0):
        pts = as_points(pts)
        self._check(self._f("map_add")(self.h, C.c_int(k), _ptr(pts), C.c_size_t(len(pts)), C.c_int(1 if fixed else 0),
                                       C.c_double(current_time)))

    def Undistort(self, pts, H0, H1, t0, t1):
        pts = as_points(pts).copy()
        self._check(self._f("undistort")(self.h, _ptr(pts), C.c_size_t(len(pts)), _ptr(_mat16(H0)), _ptr(_mat16(H1)),
                                         C.c_double(t0), C.c_double(t1)))
        return pts


_LIB = None


def lib_path():
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "_lib", "liblidarslam_b200.so")


def load_library():
    """Loads the in-tree CUDA library.  There is no CPU fallback: a missing
    library is an error, and lsb_create fails with LSB_E_NO_DEVICE without a GPU."""
    global _LIB
    if _LIB is None:
        p = lib_path()
        if not os.path.exists(p):
            raise RuntimeError("liblidarslam_b200.so is not built: run `python -c 'import __graft_entry__ as g; g.build()'`")
        _LIB = C.CDLL(p)
    return _LIB


def default_params(lib=None):
    p = Params()
    (lib or load_library()).lsb_default_params(C.byref(p))
    return p


def default_ke_params(lib=None):
    p = KeParams()
    (lib or load_library()).lsb_default_ke_params(C.byref(p))
    return p


def pcd_write(path, pts, pcd_format=PCD_BINARY_COMPRESSED, lib=None):
    """savePointCloudToPCD (reference PointCloudStorage.h:85-115): 0 ok, -1 I/O error, -3 empty cloud, -4 unknown format."""
    pts = as_points(pts)
    f = (lib or load_library()).lsb_pcd_write
    f.restype = C.c_int
    return f(os.fsencode(path), _ptr(pts), C.c_size_t(len(pts)), C.c_int(pcd_format))


def pcd_read(path, lib=None):
    """pcl::io::loadPCDFile<LidarPoint> (reference Slam.cxx:531); raises on an unreadable file."""
    lib = lib or load_library()
    f = lib.lsb_pcd_read
    f.restype = C.c_long
    n = f(os.fsencode(path), None, C.c_long(0))
    if n < 0:
        e = lib.lsb_last_error
        e.restype = C.c_char_p
        raise RuntimeError("lsb_pcd_read failed: %s" % (e() or b"").decode())
    out = np.zeros(int(n), dtype=POINT_DTYPE)
    if n:
        f(os.fsencode(path), _ptr(out), C.c_long(n))
    return out
