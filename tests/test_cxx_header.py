"""include/LidarSlam/Slam.h: the header-only C++ mirror of the reference's public interface (Slam.h:98-394) above the C ABI.
A caller written like the reference's wrappers (tests/cxx/slam_header_demo.cpp) is compiled with g++ against the header and
linked against liblidarslam_b200.so only.  CPU: it builds and fails loudly without a device.  GPU: it reproduces, bit for bit,
the trajectory the ctypes handle computes on the same frames."""
import os
import subprocess

import numpy as np
import pytest

from lidarslam_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp_path):
    exe = str(tmp_path / "slam_header_demo")
    lib_dir = os.path.dirname(capi.lib_path())
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cxx", "slam_header_demo.cpp"), "-o", exe, "-L" + lib_dir, "-llidarslam_b200",
           "-Wl,-rpath," + lib_dir]
    subprocess.check_call(cmd)
    return exe


def test_header_compiles_and_fails_loudly_without_a_device(built, tmp_path):
    exe = _build(tmp_path)
    if capi.load_library().lsb_device_count() > 0:
        pytest.skip("a CUDA device is present")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU fallback" in r.stdout


def test_header_covers_the_reference_interface():
    """Every public method of the reference class that belongs to the path is declared by the mirror (names from Slam.h:98-394)."""
    src = open(os.path.join(ROOT, "include", "LidarSlam", "Slam.h")).read()
    names = ["Reset", "AddFrame", "AddFrames", "GetWorldTransform", "GetLatencyCompensatedWorldTransform", "GetTransformCovariance",
             "GetTrajectory", "GetMap", "GetTargetSubMap", "GetKeypoints", "GetRegisteredFrame", "GetNbrFrameProcessed", "GetDebugInformation",
             "GetDebugArray", "SetWorldTransformFromGuess", "SaveMapsToPCD", "LoadMapsFromPCD", "SetUseBlobs", "GetUseBlobs",
             "GetKeyPointsExtractors", "SetKeyPointsExtractors", "GetKeyPointsExtractor", "SetKeyPointsExtractor", "GetBaseToLidarOffset",
             "SetBaseToLidarOffset", "SetWheelOdomWeight", "SetGravityWeight", "SetSensorTimeOffset", "AddGravityMeasurement",
             "AddWheelOdomMeasurement", "ClearSensorMeasurements", "SetVoxelGridDecayingThreshold", "SetVoxelGridSamplingMode", "ClearMaps",
             "SetVoxelGridLeafSize", "SetVoxelGridSize", "SetVoxelGridResolution", "SetVoxelGridMinFramesPerVoxel", "GetOverlapEstimation",
             "GetTotalMatchedKeypoints", "GetComplyMotionLimits", "SetAccelerationLimits", "SetVelocityLimits"]
    macro = ["NbThreads", "Verbosity", "EgoMotion", "Undistortion", "LoggingTimeout", "TwoDMode", "EgoMotionLMMaxIter", "EgoMotionICPMaxIter",
             "EgoMotionMaxNeighborsDistance", "EgoMotionEdgeNbNeighbors", "EgoMotionEdgeMinNbNeighbors", "EgoMotionPlaneNbNeighbors",
             "EgoMotionPlanarityThreshold", "EgoMotionEdgeMaxModelError", "EgoMotionPlaneMaxModelError", "EgoMotionInitSaturationDistance",
             "EgoMotionFinalSaturationDistance", "LocalizationLMMaxIter", "LocalizationICPMaxIter", "LocalizationMaxNeighborsDistance",
             "LocalizationEdgeNbNeighbors", "LocalizationEdgeMinNbNeighbors", "LocalizationPlaneNbNeighbors", "LocalizationPlanarityThreshold",
             "LocalizationEdgeMaxModelError", "LocalizationPlaneMaxModelError", "LocalizationBlobNbNeighbors",
             "LocalizationInitSaturationDistance", "LocalizationFinalSaturationDistance", "KfDistanceThreshold", "KfAngleThreshold", "MapUpdate",
             "OverlapSamplingRatio", "TimeWindowDuration"]
    missing = [n for n in names if (" " + n + "(") not in src] + [m for m in macro if ("LSB_PARAM(" + m + ",") not in src]
    assert not missing, missing


@pytest.mark.gpu
def test_header_caller_reproduces_the_ctypes_trajectory(gpu, synth, tmp_path):
    exe = _build(tmp_path)
    seq, sens = synth.Sequence(scene_seed=1234), synth.sensor("vlp16")
    n = 50
    g = gpu()
    p = g.GetParams()
    p.use_keypoints[2] = 1
    p.overlap_sampling_ratio = 0.33
    g.SetParams(p)
    path = str(tmp_path / "frames.bin")
    want = []
    with open(path, "wb") as f:
        for i in range(n):
            pts, _, stamp = seq.frame(sens, i)
            f.write(np.uint64(stamp).tobytes())
            f.write(pts.tobytes())
            rc = g.AddFrame(pts, stamp, i + 1)
            T, t = g.GetWorldTransform()
            want.append((rc, len(g.GetKeypoints(0)), len(g.GetKeypoints(1)), len(g.GetKeypoints(2)), T, t))
    r = subprocess.run([exe, path, str(n), str(len(pts)), str(tmp_path / "cxx_")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.strip().splitlines()
    assert len(lines) == n + 2
    az = lines.pop().split()
    assert az[0] == "azimuth" and int(az[3]) == 1                                     # estimated once, on the first frame
    assert float.fromhex(az[1]) == float(g.GetExtractorParams().azimuthal_resolution) > 0
    for i, line in enumerate(lines[:n]):
        tok = line.split()
        rc, ne, np_, nb = (int(v) for v in tok[1:5])
        T = np.array([float.fromhex(v) for v in tok[5:21]]).reshape(4, 4)
        assert (rc, ne, np_, nb) == want[i][:4], (i, line)
        assert np.array_equal(T, want[i][4]) and float.fromhex(tok[21]) == want[i][5]
    last = lines[n].split()
    assert float.fromhex(last[1]) == float(np.float32(g.GetOverlapEstimation())) and int(last[-1]) == n
    assert [int(v) for v in last[5:8]] == [len(g.GetMap(k)) for k in range(3)]
    # the maps it saved decode to the maps of the ctypes handle
    for k, kind in enumerate(("edges", "planes", "blobs")):
        assert capi.pcd_read(str(tmp_path / ("cxx_" + kind + ".pcd"))).tobytes() == g.GetMap(k, True).tobytes()
    assert "Localization skipped" in r.stderr                 # first frame: the reference's own message on stderr (Slam.cxx:1098-1107)


def test_helper_thread_class(built, tmp_path):
    """lidarslam_b200/csrc/host/worker.h (deferred map update, staged inputs): FIFO order, wait / done / drain, exceptions
    reach the waiter, sleeping helpers wake up - host logic, no GPU (tests/cxx/worker_test.cpp)."""
    exe = str(tmp_path / "worker_test")
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(cuda, "include"),
                           os.path.join(ROOT, "tests", "cxx", "worker_test.cpp"), "-o", exe, "-L" + os.path.join(cuda, "lib64"), "-lcudart",
                           "-lpthread", "-Wl,-rpath," + os.path.join(cuda, "lib64")])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "worker ok" in r.stdout, r.stdout + r.stderr
