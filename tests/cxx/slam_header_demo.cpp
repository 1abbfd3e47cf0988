// Test program for include/LidarSlam/Slam.h: a caller written against the reference's public
// interface (what LidarSlamNode::ScanCallback / vtkSlam::RequestData do: configure, AddFrame, read
// the pose, the keypoints and the maps), linked against liblidarslam_b200.so only.
//   slam_header_demo <frames.bin> <n_frames> <n_points> [pcd_prefix]
// frames.bin: per frame a uint64 stamp (us) followed by n_points LidarPoints (32 B each).
// Prints one line per frame: status, keypoint counts, the 16 entries of the world transform (hex floats).
#include <LidarSlam/Slam.h>
#include <cstdio>
#include <cstdlib>

int main(int argc, char** argv)
{
  std::unique_ptr<LidarSlam::Slam> slam;
  try { slam.reset(new LidarSlam::Slam()); }
  catch (const std::exception& e) { std::printf("no device: %s\n", e.what()); return 3; }
  if (argc < 4) { std::printf("usage: %s frames.bin n_frames n_points [pcd_prefix]\n", argv[0]); return 2; }
  const int nFrames = std::atoi(argv[2]);
  const std::size_t nPoints = (std::size_t)std::atol(argv[3]);
  std::FILE* f = std::fopen(argv[1], "rb");
  if (!f) return 2;

  // the calls a wrapper makes at start-up (LidarSlamNode.cxx:690-880)
  slam->SetUseBlobs(true);
  slam->SetUndistortion(LidarSlam::UndistortionMode::REFINED);
  slam->SetEgoMotion(LidarSlam::EgoMotionMode::MOTION_EXTRAPOLATION);
  slam->SetLocalizationICPMaxIter(3);
  slam->SetOverlapSamplingRatio(0.33f);
  auto ke = std::make_shared<LidarSlam::SpinningSensorKeypointExtractor>();
  ke->SetMinDistanceToSensor(1.5f);
  slam->SetKeyPointsExtractor(ke, 0);
  slam->SetBaseToLidarOffset(LidarSlam::Isometry3d::Identity(), 0);
  slam->SetVoxelGridLeafSize(LidarSlam::EDGE, 0.30);

  // frames are read one ahead; from the middle of the run on the next one is handed in before the current one is added
  // (PrefetchFrames: same results, its extraction overlaps the current frame)
  auto readFrame = [&](int i) -> LidarSlam::Slam::PointCloud::Ptr {
    LidarSlam::Slam::PointCloud::Ptr pc(new LidarSlam::Slam::PointCloud);
    std::uint64_t stamp = 0;
    if (std::fread(&stamp, 8, 1, f) != 1) return nullptr;
    pc->resize(nPoints);
    if (std::fread(pc->points.data(), sizeof(LidarSlam::LidarPoint), nPoints, f) != nPoints) return nullptr;
    pc->header.stamp = stamp;
    pc->header.seq = (std::uint32_t)(i + 1);
    pc->header.frame_id = "lidar";
    return pc;
  };
  LidarSlam::Slam::PointCloud::Ptr next = readFrame(0);
  for (int i = 0; i < nFrames; ++i) {
    LidarSlam::Slam::PointCloud::Ptr pc = next;
    if (!pc) return 2;
    next = (i + 1 < nFrames) ? readFrame(i + 1) : nullptr;
    if (next && i >= nFrames / 2) slam->PrefetchFrames({next});
    slam->AddFrame(pc);
    const LidarSlam::Transform T = slam->GetWorldTransform();
    std::printf("%d %d %zu %zu %zu", i, slam->GetLastStatus(), slam->GetKeypoints(LidarSlam::EDGE)->size(),
                slam->GetKeypoints(LidarSlam::PLANE)->size(), slam->GetKeypoints(LidarSlam::BLOB, true)->size());
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) std::printf(" %a", T.transform(r, c));
    std::printf(" %a\n", T.time);
  }
  const auto info = slam->GetDebugInformation();
  std::printf("overlap %a planes_used %.0f maps %zu %zu %zu frames %u\n", (double)slam->GetOverlapEstimation(), info.at("Localization: planes used"),
              slam->GetMap(LidarSlam::EDGE)->size(), slam->GetMap(LidarSlam::PLANE)->size(), slam->GetMap(LidarSlam::BLOB)->size(),
              slam->GetNbrFrameProcessed());
  // the estimate made on the first frame lives in the extractor object from then on (SSKE.cxx:169)
  std::printf("azimuth %a estimations %u\n", (double)ke->GetAzimuthalResolution(), lsb_get_azimuth_estimations(slam->GetHandle(), 0));
  if (argc > 4) slam->SaveMapsToPCD(argv[4], LidarSlam::PCDFormat::BINARY_COMPRESSED, true);
  std::fclose(f);
  return 0;
}
