// Kernel-level entry points of the C ABI: each runs ONE stage of the path on
// caller-supplied HOST data through the same kernels lsb_add_frames uses
// (parity tests and benchmarks).
#include "slam.cuh"
#include <cstring>
#include <string>

using namespace lsb;

// shares the thread-local error message of capi.cu
extern "C" void lsb_internal_set_error(const char* m);

template <typename F>
static int guarded(F&& f)
{
  try { return f(); }
  catch (const CudaError& e) { lsb_internal_set_error(e.what()); return e.code; }
  catch (const std::exception& e) { lsb_internal_set_error(e.what()); return LSB_E_INVALID; }
}

static Iso3d FromRowMajor(const double* T)
{
  Iso3d r;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) r.R(i, j) = T[i * 4 + j];
  r.t = Vec3d(T[3], T[7], T[11]);
  return r;
}
static void ToRowMajor(const Iso3d& a, double* T)
{
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) T[i * 4 + j] = a.R(i, j);
    T[i * 4 + 3] = a.t[i];
  }
  T[12] = T[13] = T[14] = 0.; T[15] = 1.;
}

static void Upload(lsb_slam* h, DevBuf<lsb_point>& buf, const lsb_point* src, size_t n)
{
  buf.reserve(std::max<size_t>(n, 1), h->stream);
  if (n) LSB_CUDA(cudaMemcpyAsync(buf.p, src, n * sizeof(lsb_point), cudaMemcpyHostToDevice, h->stream));
}

extern "C" {

int lsb_map_load_points(lsb_slam* h, int type, const lsb_point* pts, long n, int fixed)
{
  if (type < 0 || type > 2 || n < 0) return LSB_E_INVALID;
  return guarded([&] {
    h->Quiesce(false);
    Upload(h, h->tmp_cloud, pts, (size_t)n);
    h->LocalMaps[type].Add(h->tmp_cloud.p, (size_t)n, fixed != 0, 0., true, h->stream, h->lc);   // Slam.cxx:537
    LSB_CUDA(cudaStreamSynchronize(h->stream));
    return LSB_OK;
  });
}

static lsb_conversion_params ConversionDefaults(const lsb_conversion_params* p)
{
  lsb_conversion_params d;
  d.laser_id_mapping = nullptr; d.n_laser_id_mapping = 0; d.device_id = 0; d.rpm = 600.; d.timestamp_first_packet = 0;   // VelodyneToLidarNode.h:69-75
  return p ? *p : d;
}
static long ConvertedOut(lsb_slam* h, lsb_point* out, long cap)
{
  const size_t n = h->conv.n_out;
  if (out && n && cap > 0) LSB_CUDA(cudaMemcpyAsync(out, h->conv.out.p, std::min<size_t>(n, (size_t)cap) * sizeof(lsb_point), cudaMemcpyDeviceToHost, h->stream));
  LSB_CUDA(cudaStreamSynchronize(h->stream));
  return (long)n;
}
long lsb_convert_velodyne(lsb_slam* h, const lsb_velodyne_point* pts, size_t n, const lsb_conversion_params* p, lsb_point* out, long cap)
{
  if (n && !pts) return LSB_E_INVALID;
  const lsb_conversion_params cp = ConversionDefaults(p);
  if (!(cp.rpm > 0.)) return LSB_E_INVALID;
  long r = 0;
  const int rc = guarded([&] {
    h->Quiesce(false);
    h->conv.velodyne(pts, n, cp, h->stream, h->lc);
    r = ConvertedOut(h, out, cap);
    return LSB_OK;
  });
  return rc < 0 ? rc : r;
}
long lsb_convert_robosense(lsb_slam* h, const lsb_xyzi_point* pts, uint32_t width, uint32_t height, const lsb_conversion_params* p, lsb_point* out, long cap)
{
  if ((size_t)width * height && !pts) return LSB_E_INVALID;
  const lsb_conversion_params cp = ConversionDefaults(p);
  if (!(cp.rpm > 0.)) return LSB_E_INVALID;
  long r = 0;
  const int rc = guarded([&] {
    h->Quiesce(false);
    h->conv.robosense(pts, width, height, cp, h->stream, h->lc);
    r = ConvertedOut(h, out, cap);
    return LSB_OK;
  });
  return rc < 0 ? rc : r;
}
int lsb_add_converted_frame(lsb_slam* h, uint64_t stamp_us, uint32_t seq)
{
  return guarded([&] {
    h->Quiesce(true);
    const lsb_point* frame = h->conv.out.p;
    const size_t n = h->conv.n_out;
    return h->AddFrames(1, &frame, &n, &stamp_us, &seq, true);
  });
}

int lsb_sort_pairs32(lsb_slam* h, uint32_t* keys, uint32_t* values, size_t n, int key_bits)
{
  if (n && (!keys || !values)) return LSB_E_INVALID;
  return guarded([&] {
    h->Quiesce(false);
    if (n == 0) return (int)LSB_OK;
    DevBuf<uint32_t> ka, va, kb, vb;
    ka.reserve(n, h->stream); va.reserve(n, h->stream); kb.reserve(n, h->stream); vb.reserve(n, h->stream);
    LSB_CUDA(cudaMemcpyAsync(ka.p, keys, n * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
    LSB_CUDA(cudaMemcpyAsync(va.p, values, n * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
    const bool in_b = h->LocalMaps[0].sorter.sort(ka.p, va.p, kb.p, vb.p, n, key_bits, h->stream, h->lc);
    LSB_CUDA(cudaMemcpyAsync(keys, in_b ? kb.p : ka.p, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
    LSB_CUDA(cudaMemcpyAsync(values, in_b ? vb.p : va.p, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
    LSB_CUDA(cudaStreamSynchronize(h->stream));
    return (int)LSB_OK;
  });
}

int lsb_extract_keypoints(lsb_slam* h, const lsb_point* pts, size_t n, uint8_t* label_out, uint32_t counts[3])
{
  return guarded([&] {
    h->Quiesce(false);
    Upload(h, h->tmp_cloud, pts, n);
    ExtractorGpu& ke = h->Extractor(0);
    ke.compute(h->tmp_cloud.p, n, h->stream, h->lc);
    for (int k = 0; k < 3; ++k) counts[k] = ke.counts[k];
    if (label_out && n) {
      h->tmp_u8.reserve(n, h->stream);
      LSB_CUDA(cudaMemsetAsync(h->tmp_u8.p, 0, n, h->stream));
      ke.labels_input_order(h->tmp_u8.p, h->stream, h->lc);
      LSB_CUDA(cudaMemcpyAsync(label_out, h->tmp_u8.p, n, cudaMemcpyDeviceToHost, h->stream));
    }
    LSB_CUDA(cudaStreamSynchronize(h->stream));
    return LSB_OK;
  });
}

long lsb_get_extracted_keypoints(lsb_slam* h, int type, lsb_point* out, long cap)
{
  if (type < 0 || type > 2) return LSB_E_INVALID;
  return guarded([&] {
    h->Quiesce(false);
    ExtractorGpu& ke = h->Extractor(0);
    const size_t n = ke.counts[type];
    if (out && n && cap > 0) {
      LSB_CUDA(cudaMemcpyAsync(out, ke.kp[type].p, std::min<size_t>(n, (size_t)cap) * sizeof(lsb_point), cudaMemcpyDeviceToHost, h->stream));
      LSB_CUDA(cudaStreamSynchronize(h->stream));
    }
    return (int)n;
  });
}

int lsb_knn_search(lsb_slam* h, const lsb_point* cloud, size_t m, const float* queries_xyz, size_t nq, int k, int32_t* idx_out, float* d2_out)
{
  return guarded([&] {
    h->Quiesce(false);
    if (k < 1 || k > KNN_MAX_K) throw CudaError(LSB_E_INVALID, "k must be in [1, 32]");
    Upload(h, h->tmp_cloud, cloud, m);
    KnnGrid grid;
    DevBuf<float> dq, dd2;
    DevBuf<int32_t> didx;
    dq.reserve(std::max<size_t>(3 * nq, 1), h->stream); dd2.reserve(std::max<size_t>(nq * k, 1), h->stream); didx.reserve(std::max<size_t>(nq * k, 1), h->stream);
    if (nq) LSB_CUDA(cudaMemcpyAsync(dq.p, queries_xyz, 3 * nq * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    if (m) grid.build(h->tmp_cloud.p, m, 1.0f, nullptr, nullptr, h->stream, h->lc);
    grid.search(dq.p, nq, k, -1.f, didx.p, dd2.p, nullptr, /*entry_major=*/false, h->stream, h->lc);
    if (nq) {
      LSB_CUDA(cudaMemcpyAsync(idx_out, didx.p, nq * k * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
      LSB_CUDA(cudaMemcpyAsync(d2_out, dd2.p, nq * k * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    }
    LSB_CUDA(cudaStreamSynchronize(h->stream));
    return LSB_OK;
  });
}

int lsb_measure_fp_peaks(int cuda_device, double out_tflops[4])
{
  return guarded([&] { measure_fp_peaks(cuda_device, out_tflops); return LSB_OK; });
}

// no allocation, no handle, no synchronisation: see the declaration (no reference counterpart)
int lsb_knn_search_noalloc(int cuda_device, const lsb_point* d_cloud, size_t m, const float* d_queries_xyz, size_t nq, int k, int32_t* d_idx_out,
                           float* d_d2_out, void* cuda_stream)
{
  return guarded([&] {
    knn_search_noalloc(cuda_device, d_cloud, m, d_queries_xyz, nq, k, d_idx_out, d_d2_out, static_cast<cudaStream_t>(cuda_stream));
    return LSB_OK;
  });
}

int lsb_match_and_solve(lsb_slam* h, const lsb_point* const keypoints[3], const size_t nk[3], const lsb_point* const target[3],
                        const size_t nt[3], const lsb_match_params* mp, int lm_max_iter, const double prior_rowmajor[16],
                        uint8_t* status_out, double* weight_out, double pose_out_rowmajor[16], int* lm_accepted_steps,
                        int* lm_iterations, double* final_cost, lsb_registration_error* err_out)
{
  return guarded([&] {
    h->Quiesce(false);
    const Iso3d T = FromRowMajor(prior_rowmajor);
    DevBuf<lsb_point> dk[3], dt[3];
    KnnGrid grids[3];
    for (int k = 0; k < 3; ++k) {
      Upload(h, dk[k], keypoints[k], nk[k]);
      Upload(h, dt[k], target[k], nt[k]);
      if (nt[k]) grids[k].build(dt[k].p, nt[k], 1.0f, nullptr, nullptr, h->stream, h->lc);
    }
    MatchWorkspace& mw = h->mw;
    mw.begin_iteration(nk, h->stream);
    for (int k = 0; k < 3; ++k) mw.match_type(k, dk[k].p, nk[k], grids[k], T, *mp, h->stream, h->lc);
    int hist[3][LSB_MATCH_N_STATUS];
    mw.read_histograms(hist, h->stream);
    if (mw.total) {
      LSB_CUDA(cudaMemcpyAsync(status_out, mw.status.p, mw.total, cudaMemcpyDeviceToHost, h->stream));
      LSB_CUDA(cudaMemcpyAsync(weight_out, mw.weight.p, mw.total * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
      LSB_CUDA(cudaStreamSynchronize(h->stream));
    }
    unsigned total = 0;
    for (int k = 0; k < 3; ++k) total += (unsigned)hist[k][LSB_MATCH_SUCCESS];
    if (total == 0) {
      ToRowMajor(T, pose_out_rowmajor);
      *lm_accepted_steps = 0; *lm_iterations = 0; *final_cost = 0.;
      if (err_out) std::memset(err_out, 0, sizeof(*err_out));
      return LSB_W_LOC_SKIPPED;
    }
    double x0[6];
    iso_to_xyzrpy(T, x0);
    const LmState st = mw.solve(x0, lm_max_iter, h->P.two_d_mode != 0, mp->saturation_distance, 1u, hist, h->stream, h->lc);
    ToRowMajor(xyzrpy_to_iso(st.x), pose_out_rowmajor);
    *lm_accepted_steps = st.num_successful;
    *lm_iterations = st.iter;
    *final_cost = st.cost;
    if (err_out) registration_error_from_normal_matrix(st.H, h->P.two_d_mode != 0, err_out);
    return LSB_OK;
  });
}

int lsb_map_add(lsb_slam* h, int type, const lsb_point* pts, size_t n, int fixed, double current_time)
{
  if (type < 0 || type > 2) return LSB_E_INVALID;
  return guarded([&] {
    h->Quiesce(false);
    Upload(h, h->tmp_cloud, pts, n);
    h->LocalMaps[type].Add(h->tmp_cloud.p, n, fixed != 0, current_time, true, h->stream, h->lc);
    LSB_CUDA(cudaStreamSynchronize(h->stream));
    return LSB_OK;
  });
}

int lsb_undistort(lsb_slam* h, lsb_point* pts, size_t n, const double H0[16], const double H1[16], double t0, double t1)
{
  return guarded([&] {
    h->Quiesce(false);
    Interpolator it;
    it.t0 = t0; it.t1 = t1;
    it.set_transforms(FromRowMajor(H0), FromRowMajor(H1));
    Upload(h, h->tmp_cloud, pts, n);
    dev_undistort(h->tmp_cloud.p, n, it, 0.0, h->stream, h->lc);
    if (n) LSB_CUDA(cudaMemcpyAsync(pts, h->tmp_cloud.p, n * sizeof(lsb_point), cudaMemcpyDeviceToHost, h->stream));
    LSB_CUDA(cudaStreamSynchronize(h->stream));
    return LSB_OK;
  });
}

}  // extern "C"
