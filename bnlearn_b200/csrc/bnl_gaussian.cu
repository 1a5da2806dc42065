// bnl_gaussian.cu -- host side of the Gaussian entry points.
#include <string.h>
#include <cmath>
#include <algorithm>
#include <thread>

#include "bnl_gaussian_kernels.cuh"

// Means and covariance of all numeric columns on one device; called once per
// data set from bnl_data_register().
int bnl_gaussian_prepare(bnl_ctx *ctx, BnlDataSet &ds, int di) {
  BnlDevice &dev = ctx->devs[di];
  BnlDeviceData &dd = ds.dev[di];
  const int p = ds.nc;
  BNL_CUDA(ctx, cudaSetDevice(dev.id));
  int prc;
  dd.mean_bytes = (size_t)p * sizeof(double);
  dd.cov_bytes = (size_t)p * p * sizeof(double);
  if ((prc = bnl_pool_alloc(ctx, dev, (void **)&dd.mean, dd.mean_bytes))) return prc;
  if ((prc = bnl_pool_alloc(ctx, dev, (void **)&dd.cov, dd.cov_bytes))) return prc;

  // row slices: enough CTAs for several waves, at least 8192 rows per slice,
  // and at most 512 MiB of partials.
  const int ntile = (p + GRAM_TILE - 1) / GRAM_TILE;
  const int ntri = ntile * (ntile + 1) / 2;
  long long nsplit = (12ll * dev.sm_count + ntri - 1) / ntri;
  nsplit = std::min(nsplit, std::max(1ll, ds.nobs / 8192));
  nsplit = std::min(nsplit, std::max(1ll, (512ll << 20) / ((long long)p * p * 8)));
  nsplit = std::max(nsplit, 1ll);
  long long msplit = std::min<long long>(std::max(1ll, ds.nobs / 65536), 64);

  // The Gram kernel reads X - mean from scratch memory when the whole centred data set fits
  // beside everything else (16 GB for 2000 columns x 1 M rows: 133 ms for the moments against
  // 140); otherwise it centres each stage in shared memory. (Slabs of rows through a smaller
  // scratch buffer were slower than no scratch at all -- a partial last wave per slab.)
  // BNL_GRAM_SCRATCH=0 / 1 forces the choice (tests).
  const long long rows_pad = (ds.nobs + GRAM_KB - 1) / GRAM_KB * GRAM_KB;
  bool use_scratch;
  {
    size_t free_b = 0, total_b = 0;
    cudaMemGetInfo(&free_b, &total_b);
    size_t cached = 0;
    for (auto &b : dev.pool) cached += b.second;
    const double avail = std::max(0.0, (double)free_b + (double)cached - (double)(2ull << 30));
    use_scratch = (double)p * (double)rows_pad * 8.0 <= 0.75 * avail;
    if (const char *v = getenv("BNL_GRAM_SCRATCH")) if (*v) use_scratch = atoi(v) != 0;
  }

  double *partial = nullptr, *scratch = nullptr;
  size_t pbytes = std::max((size_t)nsplit * p * p, (size_t)msplit * p) * sizeof(double);
  const size_t sbytes = use_scratch ? (size_t)rows_pad * p * sizeof(double) : 0;
  if ((prc = bnl_pool_alloc(ctx, dev, (void **)&partial, pbytes))) return prc;
  if (sbytes && (prc = bnl_pool_alloc(ctx, dev, (void **)&scratch, sbytes))) { bnl_pool_free(dev, partial, pbytes); return prc; }
  cudaError_t e;

  cudaEventRecord(dev.evc0, dev.stream);
  col_sum_kernel<<<dim3(p, (unsigned)msplit), 256, 0, dev.stream>>>(dd.f64, ds.stride_f64, ds.nobs, p, partial);
  col_mean_kernel<<<(p + 127) / 128, 128, 0, dev.stream>>>(partial, (int)msplit, p, ds.nobs, dd.mean);
  int launches = 4;
  if (scratch) {
    const unsigned gx = (unsigned)std::min<long long>((rows_pad / 2 + 255) / 256, 4096);
    gram_centre_kernel<<<dim3(gx, p), 256, 0, dev.stream>>>(dd.f64, ds.stride_f64, ds.nobs, rows_pad, dd.mean, scratch);
    cudaFuncSetAttribute(gram_dmma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, GRAM_SMEM);
    gram_dmma_kernel<true><<<dim3(ntri, (unsigned)nsplit), GRAM_THREADS, GRAM_SMEM, dev.stream>>>(
        scratch, (size_t)rows_pad, rows_pad, p, dd.mean, partial, ntile);
    launches = 5;
  } else {
    cudaFuncSetAttribute(gram_dmma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, GRAM_SMEM);
    gram_dmma_kernel<false><<<dim3(ntri, (unsigned)nsplit), GRAM_THREADS, GRAM_SMEM, dev.stream>>>(
        dd.f64, ds.stride_f64, ds.nobs, p, dd.mean, partial, ntile);
  }
  const long long elems = (long long)p * p;
  gram_finish_kernel<<<(unsigned)((elems + 255) / 256), 256, 0, dev.stream>>>(partial, (int)nsplit, p,
                                                                              ds.nobs, dd.cov);
  cudaEventRecord(dev.evc1, dev.stream);
  e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaStreamSynchronize(dev.stream);
  bnl_pool_free(dev, partial, pbytes);
  // a large scratch block goes straight back to the driver: cached in the pool it would hold
  // as much device memory as the data set itself until something evicted it
  if (scratch && sbytes > ((size_t)1 << 30)) cudaFree(scratch);
  else if (scratch) bnl_pool_free(dev, scratch, sbytes);
  if (e == cudaSuccess && di == 0) {
    // bnl_last_call_profile() after a registration reports the moments kernels.
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, dev.evc0, dev.evc1) == cudaSuccess) {
      ctx->last_ms = ctx->last_count_ms = ms;
      ctx->last_epi_ms = 0.0;
      ctx->last_launches = launches;
    }
  }
  if (e != cudaSuccess)
    return bnl_fail(ctx, BNL_ECUDA, "Gaussian moments failed: %s", cudaGetErrorString(e));
  if (di == 0) {
    ds.h_mean.resize(p);
    ds.h_cov.resize((size_t)p * p);
    BNL_CUDA(ctx, cudaMemcpy(ds.h_mean.data(), dd.mean, (size_t)p * sizeof(double), cudaMemcpyDeviceToHost));
    BNL_CUDA(ctx, cudaMemcpy(ds.h_cov.data(), dd.cov, (size_t)p * p * sizeof(double), cudaMemcpyDeviceToHost));
    // a column with a missing value has a NaN mean: such data sets take the
    // per-unit complete-row moments instead of the global matrix (run_gblock).
    ds.any_nan = false;
    for (int c = 0; c < p; c++) if (std::isnan(ds.h_mean[c])) ds.any_nan = true;
  }
  return BNL_OK;
}

namespace {

struct GRequest {
  int kind;              // 0 tests, 1 scores
  int code;
  double arg;
  int n, kmax;
  const int *a, *b;      // x[] and y[] (tests) or target[] (scores; b unused)
  const int *off, *idx;  // CSR conditioning sets / parents
  double *o0, *o1, *o2;
};

// Units [b, e) on device slot d: the caller's arrays go to the device as they
// are (no per-unit descriptor is built on the host).
int run_gblock(bnl_ctx *ctx, BnlDataSet &ds, int d, GRequest &rq, int b, int e, BnlTiming *tm) {
  BnlDevice &dev = ctx->devs[d];
  BnlDeviceData &dd = ds.dev[d];
  BNL_CUDA(ctx, cudaSetDevice(dev.id));
  // incomplete data: per-unit moments need scratch proportional to units x row
  // chunks, so long requests are cut into sub-blocks (also: grid.y <= 65535).
  const bool shrink = rq.kind == 0 && rq.code == BNL_MI_G_SH;     // needs a pass over the rows per unit
  const bool masked = ds.any_nan || shrink;
  const int m_kb = rq.kmax <= 4 ? 4 : (rq.kmax <= 8 ? 8 : 16);
  const int m_np = m_kb * (m_kb + 1) / 2;
  const int m_chunks = (int)std::max(1ll, std::min(256ll, (ds.nobs + 4095) / 4096));
  const size_t m_per_unit = ((size_t)m_chunks * (BNL_MAXK + 1 + m_np + 1) + (size_t)rq.kmax * (rq.kmax + 1) + 1) *
                            sizeof(double);
  if (masked) {
    const long long max_units = std::max(1ll, std::min(65535ll, (long long)((256ull << 20) / m_per_unit)));
    if (e - b > max_units) {
      for (int s = b; s < e; s += (int)max_units) {
        const int rc = run_gblock(ctx, ds, d, rq, s, std::min(e, s + (int)max_units), tm);
        if (rc) return rc;
      }
      return BNL_OK;
    }
  }
  if (!dev.gattr_done) {
    BNL_CUDA(ctx, cudaFuncSetAttribute(gauss_ci_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 1024));
    BNL_CUDA(ctx, cudaFuncSetAttribute(gauss_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dev.max_smem - 1024));
    dev.gattr_done = true;
  }
  const int cnt = e - b;
  const int i0 = rq.off[b], nidx = rq.off[e] - i0;
  // device layout of the staged inputs: a | b | off (cnt + 1) | idx | slot map | err
  const size_t n_a = cnt, n_b = rq.kind == 0 ? cnt : 0, n_off = (size_t)cnt + 1, n_idx = nidx;
  const size_t total = n_a + n_b + n_off + n_idx + ds.ncols + 1;
  int rc;
  if ((rc = bnl_reserve_device(ctx, &dev.d_gidx, &dev.gidx_cap, total))) return rc;
  if ((rc = bnl_reserve_pinned(ctx, &dev.h_gidx, &dev.h_gidx_cap, n_off + ds.ncols + 1))) return rc;
  if ((rc = bnl_reserve_device(ctx, &dev.d_out, &dev.out_cap, (size_t)cnt * 4))) return rc;
  if ((rc = bnl_reserve_pinned(ctx, &dev.h_out, &dev.h_out_cap, (size_t)cnt * 4))) return rc;
  int *d_a = dev.d_gidx, *d_b = d_a + n_a, *d_off = d_b + n_b, *d_idx = d_off + n_off;
  int *d_slot = d_idx + n_idx, *d_err = d_slot + ds.ncols;
  // offsets rebased to this block; slot map; cleared error flag.
  for (int i = 0; i <= cnt; i++) dev.h_gidx[i] = rq.off[b + i] - i0;
  for (int c = 0; c < ds.ncols; c++) dev.h_gidx[n_off + c] = ds.is_numeric[c] ? ds.slot[c] : -1;
  dev.h_gidx[n_off + ds.ncols] = 0;

  const int mats = rq.kind == 0 ? 2 : 1;
  const size_t per_thread = (size_t)mats * rq.kmax * rq.kmax * sizeof(double);
  int tpb = 128;
  while (tpb > 32 && per_thread * tpb > (size_t)dev.max_smem - 2048) tpb >>= 1;
  const size_t smem = per_thread * tpb;
  if (smem > (size_t)dev.max_smem - 1024)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "unit with %d columns needs too much shared memory", rq.kmax);

  BNL_CUDA(ctx, cudaEventRecord(dev.ev0, dev.stream));
  BNL_CUDA(ctx, cudaMemcpyAsync(d_a, rq.a + b, n_a * sizeof(int), cudaMemcpyHostToDevice, dev.stream));
  if (n_b)
    BNL_CUDA(ctx, cudaMemcpyAsync(d_b, rq.b + b, n_b * sizeof(int), cudaMemcpyHostToDevice, dev.stream));
  BNL_CUDA(ctx, cudaMemcpyAsync(d_off, dev.h_gidx, n_off * sizeof(int), cudaMemcpyHostToDevice, dev.stream));
  if (n_idx)
    BNL_CUDA(ctx, cudaMemcpyAsync(d_idx, rq.idx + i0, n_idx * sizeof(int), cudaMemcpyHostToDevice, dev.stream));
  BNL_CUDA(ctx, cudaMemcpyAsync(d_slot, dev.h_gidx + n_off, (size_t)(ds.ncols + 1) * sizeof(int),
                                cudaMemcpyHostToDevice, dev.stream));
  const int grid = (cnt + tpb - 1) / tpb;
  BNL_CUDA(ctx, cudaEventRecord(dev.evc0, dev.stream));
  double *mblock = nullptr, *ucov = nullptr, *umean = nullptr, *ushr = nullptr;
  long long *unobs = nullptr;
  const size_t mbytes = m_per_unit * (size_t)cnt;
  if (masked) {
    if ((rc = bnl_pool_alloc(ctx, dev, (void **)&mblock, mbytes))) return rc;
    double *part1 = mblock, *part2 = part1 + (size_t)cnt * m_chunks * (BNL_MAXK + 1);
    umean = part2 + (size_t)cnt * m_chunks * m_np;
    ucov = umean + (size_t)cnt * rq.kmax;
    unobs = reinterpret_cast<long long *>(ucov + (size_t)cnt * rq.kmax * rq.kmax);
    ushr = reinterpret_cast<double *>(unobs + cnt);
    const long long chunk_rows = (ds.nobs + m_chunks - 1) / m_chunks;
    const int *d_b_or_null = rq.kind == 0 ? d_b : nullptr;
    const dim3 mg((unsigned)m_chunks, (unsigned)cnt);
    masked_sums_kernel<<<mg, BNL_MASK_THREADS, 0, dev.stream>>>(dd.f64, ds.stride_f64, ds.nobs, chunk_rows, d_a,
                                                                d_b_or_null, d_off, d_idx, d_slot, ds.ncols, part1);
    masked_means_kernel<<<(cnt * (rq.kmax + 1) + 127) / 128, 128, 0, dev.stream>>>(part1, m_chunks, cnt, rq.kmax,
                                                                                   umean, unobs);
    if (m_kb == 4)
      masked_cross_kernel<4><<<mg, BNL_MASK_THREADS, 0, dev.stream>>>(dd.f64, ds.stride_f64, ds.nobs, chunk_rows,
          d_a, d_b_or_null, d_off, d_idx, d_slot, ds.ncols, rq.kmax, umean, part2);
    else if (m_kb == 8)
      masked_cross_kernel<8><<<mg, BNL_MASK_THREADS, 0, dev.stream>>>(dd.f64, ds.stride_f64, ds.nobs, chunk_rows,
          d_a, d_b_or_null, d_off, d_idx, d_slot, ds.ncols, rq.kmax, umean, part2);
    else
      masked_cross_kernel<16><<<mg, BNL_MASK_THREADS, 0, dev.stream>>>(dd.f64, ds.stride_f64, ds.nobs, chunk_rows,
          d_a, d_b_or_null, d_off, d_idx, d_slot, ds.ncols, rq.kmax, umean, part2);
    masked_cov_kernel<<<(cnt * m_np + 127) / 128, 128, 0, dev.stream>>>(part2, m_chunks, cnt, rq.kmax, m_kb, unobs,
                                                                        ucov);
    tm->launches += 4;
    if (shrink) {
      if (m_kb == 4)
        masked_shrink_kernel<4><<<mg, BNL_MASK_THREADS, 0, dev.stream>>>(dd.f64, ds.stride_f64, ds.nobs, chunk_rows,
            d_a, d_b_or_null, d_off, d_idx, d_slot, ds.ncols, rq.kmax, umean, ucov, unobs, ushr);
      else if (m_kb == 8)
        masked_shrink_kernel<8><<<mg, BNL_MASK_THREADS, 0, dev.stream>>>(dd.f64, ds.stride_f64, ds.nobs, chunk_rows,
            d_a, d_b_or_null, d_off, d_idx, d_slot, ds.ncols, rq.kmax, umean, ucov, unobs, ushr);
      else
        masked_shrink_kernel<16><<<mg, BNL_MASK_THREADS, 0, dev.stream>>>(dd.f64, ds.stride_f64, ds.nobs, chunk_rows,
            d_a, d_b_or_null, d_off, d_idx, d_slot, ds.ncols, rq.kmax, umean, ucov, unobs, ushr);
      tm->launches++;
    }
  }
  if (rq.kind == 0)
    gauss_ci_kernel<<<grid, tpb, smem, dev.stream>>>(d_a, d_b, d_off, d_idx, d_slot, ds.ncols, cnt, dd.cov, ds.nc,
                                                     ds.gauss_nobs > 0 ? ds.gauss_nobs : ds.nobs, rq.code, rq.kmax, dev.d_out, d_err, ucov, unobs,
                                                     ushr, m_chunks);
  else
    gauss_score_kernel<<<grid, tpb, smem, dev.stream>>>(d_a, d_off, d_idx, d_slot, ds.ncols, cnt, dd.cov, dd.mean,
                                                        ds.nc, ds.gauss_nobs > 0 ? ds.gauss_nobs : ds.nobs, rq.code,
                                                        rq.code == BNL_BGE ? ds.bge_mu : rq.arg, rq.kmax,
                                                        dev.d_out, d_err, ucov, umean, unobs,
                                                        ds.bge_w > 0.0 ? ds.bge_w : (double)(ds.ncols + 2),
                                                        dd.bge_nu ? dd.bge_nu : dd.mean);
  cudaError_t launch_err = cudaGetLastError();
  if (launch_err != cudaSuccess) {
    cudaStreamSynchronize(dev.stream);
    bnl_pool_free(dev, mblock, mbytes);
    return bnl_fail(ctx, BNL_ECUDA, "Gaussian kernels failed to launch: %s", cudaGetErrorString(launch_err));
  }
  tm->launches++;
  BNL_CUDA(ctx, cudaEventRecord(dev.evc1, dev.stream));
  BNL_CUDA(ctx, cudaMemcpyAsync(dev.h_out, dev.d_out, (size_t)cnt * 4 * sizeof(double),
                                cudaMemcpyDeviceToHost, dev.stream));
  int err = 0;
  BNL_CUDA(ctx, cudaMemcpyAsync(&err, d_err, sizeof(int), cudaMemcpyDeviceToHost, dev.stream));
  BNL_CUDA(ctx, cudaEventRecord(dev.ev1, dev.stream));
  const cudaError_t sync_err = cudaStreamSynchronize(dev.stream);
  bnl_pool_free(dev, mblock, mbytes);
  if (sync_err != cudaSuccess)
    return bnl_fail(ctx, BNL_ECUDA, "Gaussian kernels failed: %s", cudaGetErrorString(sync_err));
  if (err)
    return bnl_fail(ctx, BNL_EINVAL, "unit %d refers to a column that is out of range or not numeric",
                    b + err - 1);
  for (int i = 0; i < cnt; i++) {
    if (rq.o0) rq.o0[b + i] = dev.h_out[4 * i + 0];
    if (rq.o1) rq.o1[b + i] = dev.h_out[4 * i + 1];
    if (rq.o2) rq.o2[b + i] = dev.h_out[4 * i + 2];
  }
  BNL_CUDA(ctx, cudaEventElapsedTime(&tm->total, dev.ev0, dev.ev1));
  BNL_CUDA(ctx, cudaEventElapsedTime(&tm->count, dev.evc0, dev.evc1));
  return BNL_OK;
}

int run_grequest(bnl_ctx *ctx, BnlDataSet &ds, GRequest &rq) {
  const int ndev = (int)ctx->devs.size();
  ctx->last_ms = ctx->last_count_ms = ctx->last_epi_ms = 0.0;
  ctx->last_launches = 0;
  if (rq.n == 0) return BNL_OK;
  // widest unit (sizes the per-thread work matrices) and a sanity check of the CSR.
  const int extra = rq.kind == 0 ? 2 : 1;
  rq.kmax = extra;
  for (int u = 0; u < rq.n; u++) {
    const int k = rq.off[u + 1] - rq.off[u];
    if (k < 0) return bnl_fail(ctx, BNL_EINVAL, "offsets must be non-decreasing (unit %d)", u);
    if (k + extra > BNL_MAXK)
      return bnl_fail(ctx, BNL_EUNSUPPORTED, "a unit may involve at most %d columns (got %d)", BNL_MAXK, k + extra);
    rq.kmax = std::max(rq.kmax, k + extra);
  }
  if (rq.off[rq.n] > 0 && !rq.idx) return bnl_fail(ctx, BNL_EINVAL, "index array missing");
  std::vector<int> rcs(ndev, BNL_OK);
  std::vector<BnlTiming> tm(ndev);
  if (ndev == 1) {
    rcs[0] = run_gblock(ctx, ds, 0, rq, 0, rq.n, &tm[0]);
  } else {
    std::vector<std::thread> workers;
    for (int d = 0; d < ndev; d++) {
      int b, e;
      bnl_block(rq.n, ndev, d, &b, &e);
      if (b >= e) continue;
      workers.emplace_back([&, d, b, e]() { rcs[d] = run_gblock(ctx, ds, d, rq, b, e, &tm[d]); });
    }
    for (auto &w : workers) w.join();
    cudaSetDevice(ctx->devs[0].id);
  }
  for (int d = 0; d < ndev; d++) {
    ctx->last_ms = std::max(ctx->last_ms, (double)tm[d].total);
    ctx->last_count_ms = std::max(ctx->last_count_ms, (double)tm[d].count);
    ctx->last_launches += tm[d].launches;
  }
  for (int d = 0; d < ndev; d++)
    if (rcs[d] != BNL_OK) return rcs[d];
  return BNL_OK;
}

}  // namespace

extern "C" int bnl_ci_gaussian(bnl_ctx *ctx, bnl_data data, int test_e, int n, const int *x,
                               const int *y, const int *z_off, const int *z_idx,
                               double *stat, double *df, double *pvalue) {
  if (!ctx) return BNL_EINVAL;
  auto it = ctx->data.find(data);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  BnlDataSet &ds = it->second;
  if (test_e != BNL_COR && test_e != BNL_ZF && test_e != BNL_MI_G && test_e != BNL_MI_G_SH)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "test %d is not a Gaussian test handled here", test_e);
  return bnl_ci_gaussian_any(ctx, data, test_e, n, x, y, z_off, z_idx, stat, df, pvalue);
}

// Same request with the library-internal statistic codes allowed (BNL_MI_G_COEF).
int bnl_ci_gaussian_any(bnl_ctx *ctx, bnl_data data, int test_e, int n, const int *x, const int *y,
                        const int *z_off, const int *z_idx, double *stat, double *df,
                        double *pvalue) {
  auto it = ctx->data.find(data);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  BnlDataSet &ds = it->second;
  if (n < 0 || (n > 0 && (!x || !y || !z_off)))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_ci_gaussian: bad arguments");
  if (ds.nc == 0) return bnl_fail(ctx, BNL_EINVAL, "the data set has no numeric columns");
  // mi-g-sh needs a pass over the rows (the shrinkage intensity), which a rank of a
  // row-sharded run only holds in part: refuse rather than return a shard-local value.
  if (ds.gauss_nobs > 0 && test_e == BNL_MI_G_SH)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "mi-g-sh needs the rows, not pooled moments (bnl_data_set_moments)");
  GRequest rq{0, test_e, 0.0, n, 2, x, y, z_off, z_idx, stat, df, pvalue};
  return run_grequest(ctx, ds, rq);
}

extern "C" int bnl_ci_gaussian_allpairs(bnl_ctx *ctx, bnl_data data, int test_e, double alpha, int ncols,
                                        const int *cols, double *pvalue, int *keep_i, int *keep_j, long long cap,
                                        long long *nkeep, double *test_counter) {
  if (!ctx) return BNL_EINVAL;
  auto it = ctx->data.find(data);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  BnlDataSet &ds = it->second;
  if (test_e != BNL_COR && test_e != BNL_ZF && test_e != BNL_MI_G)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "bnl_ci_gaussian_allpairs handles cor, zf and mi-g");
  if (ncols < 0 || (ncols > 0 && !cols) || cap < 0 || (cap > 0 && (!keep_i || !keep_j)))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_ci_gaussian_allpairs: bad arguments");
  if (ds.any_nan)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "bnl_ci_gaussian_allpairs needs complete data (use bnl_ci_gaussian)");
  if (nkeep) *nkeep = 0;
  if (ncols < 2) return BNL_OK;
  if (ncols > 32768)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "bnl_ci_gaussian_allpairs: at most 32768 columns per call (the flag matrix is n x n)");
  BnlDevice &dev = ctx->devs[0];
  BnlDeviceData &dd = ds.dev[0];
  BNL_CUDA(ctx, cudaSetDevice(dev.id));
  const size_t n = (size_t)ncols;
  int rc;
  // one int block: slots [n] | rowcount [n] | total (2 ints) | keep_i [cap] | keep_j [cap]; flags and p-values apart
  const size_t ints = 2 * n + 4 + 2 * (size_t)cap;
  if ((rc = bnl_reserve_pinned(ctx, &dev.h_gidx, &dev.h_gidx_cap, ints))) return rc;
  if ((rc = bnl_reserve_device(ctx, &dev.d_gidx, &dev.gidx_cap, ints))) return rc;
  for (int i = 0; i < ncols; i++) {
    const int c = cols[i];
    if (c < 0 || c >= ds.ncols || !ds.is_numeric[c])
      return bnl_fail(ctx, BNL_EINVAL, "bnl_ci_gaussian_allpairs: column %d is out of range or not numeric", c);
    dev.h_gidx[i] = ds.slot[c];
  }
  unsigned char *d_keep = nullptr;
  double *d_p = nullptr;
  const size_t kbytes = n * n, pbytes = pvalue ? n * n * sizeof(double) : 0;
  if ((rc = bnl_pool_alloc(ctx, dev, (void **)&d_keep, kbytes))) return rc;
  if (pvalue && (rc = bnl_pool_alloc(ctx, dev, (void **)&d_p, pbytes))) { bnl_pool_free(dev, d_keep, kbytes); return rc; }
  int *d_slots = dev.d_gidx, *d_cnt = d_slots + n, *d_ki = d_cnt + n + 4, *d_kj = d_ki + cap;
  long long *d_total = reinterpret_cast<long long *>(d_cnt + n);      // int index 2n: 8-byte aligned
  auto fail = [&](cudaError_t e, const char *what) {
    cudaStreamSynchronize(dev.stream);
    bnl_pool_free(dev, d_keep, kbytes);
    bnl_pool_free(dev, d_p, pbytes);
    return bnl_fail(ctx, BNL_ECUDA, "%s failed: %s", what, cudaGetErrorString(e));
  };
  cudaError_t e;
  cudaEventRecord(dev.ev0, dev.stream);
  if ((e = cudaMemcpyAsync(d_slots, dev.h_gidx, n * sizeof(int), cudaMemcpyHostToDevice, dev.stream)) != cudaSuccess)
    return fail(e, "upload");
  cudaEventRecord(dev.evc0, dev.stream);
  const long long nobs = ds.gauss_nobs > 0 ? ds.gauss_nobs : ds.nobs;
  gauss_allpairs_kernel<<<ncols, 256, 0, dev.stream>>>(dd.cov, ds.nc, d_slots, ncols, nobs, test_e, alpha, d_p, d_keep,
                                                       d_cnt);
  gauss_allpairs_compact_kernel<<<(ncols + 3) / 4, 128, 0, dev.stream>>>(d_keep, d_cnt, ncols, cap, d_ki, d_kj, d_total);
  if ((e = cudaGetLastError()) != cudaSuccess) return fail(e, "launch");
  cudaEventRecord(dev.evc1, dev.stream);
  long long total = 0;
  if ((e = cudaMemcpyAsync(&total, d_total, sizeof(long long), cudaMemcpyDeviceToHost, dev.stream)) != cudaSuccess)
    return fail(e, "download");
  if ((e = cudaStreamSynchronize(dev.stream)) != cudaSuccess) return fail(e, "kernels");
  const long long got = std::min(total, cap);
  if (got > 0) {
    cudaMemcpyAsync(keep_i, d_ki, (size_t)got * sizeof(int), cudaMemcpyDeviceToHost, dev.stream);
    cudaMemcpyAsync(keep_j, d_kj, (size_t)got * sizeof(int), cudaMemcpyDeviceToHost, dev.stream);
  }
  if (pvalue) cudaMemcpyAsync(pvalue, d_p, pbytes, cudaMemcpyDeviceToHost, dev.stream);
  cudaEventRecord(dev.ev1, dev.stream);
  if ((e = cudaStreamSynchronize(dev.stream)) != cudaSuccess) return fail(e, "download");
  bnl_pool_free(dev, d_keep, kbytes);
  bnl_pool_free(dev, d_p, pbytes);
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, dev.ev0, dev.ev1) == cudaSuccess) ctx->last_ms = ms;
  if (cudaEventElapsedTime(&ms, dev.evc0, dev.evc1) == cudaSuccess) ctx->last_count_ms = ms;
  ctx->last_epi_ms = 0.0;
  ctx->last_launches = 2;
  if (nkeep) *nkeep = total;
  if (test_counter) *test_counter += (double)ncols * (double)(ncols - 1) / 2.0;
  return BNL_OK;
}

extern "C" int bnl_score_gaussian(bnl_ctx *ctx, bnl_data data, int score_e, double k, int n,
                                  const int *target, const int *pa_off, const int *pa_idx,
                                  double *score, double *nparams) {
  if (!ctx) return BNL_EINVAL;
  auto it = ctx->data.find(data);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  BnlDataSet &ds = it->second;
  if (score_e == BNL_EBIC_G) {            // per.node.score.c:155-173
    int rc = bnl_score_gaussian(ctx, data, BNL_BIC_G, k, n, target, pa_off, pa_idx, score, nparams);
    if (rc) return rc;
    const double pen = 4.0 * ds.ebic_gamma * log((double)ds.ncols);
    for (int u = 0; u < n; u++) score[u] -= pen * (double)(pa_off[u + 1] - pa_off[u]);
    return BNL_OK;
  }
  if (score_e != BNL_LOGLIK_G && score_e != BNL_AIC_G && score_e != BNL_BIC_G && score_e != BNL_NAL_G &&
      score_e != BNL_PNAL_G && score_e != BNL_BGE)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "score %d is not a Gaussian score handled here", score_e);
  if (score_e == BNL_BGE && ds.any_nan)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "bge needs complete data (as in the reference's front end)");
  if (n < 0 || (n > 0 && (!target || !pa_off || !score)))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_score_gaussian: bad arguments");
  if (ds.nc == 0) return bnl_fail(ctx, BNL_EINVAL, "the data set has no numeric columns");
  GRequest rq{1, score_e, k, n, 1, target, nullptr, pa_off, pa_idx, score, nparams, nullptr};
  return run_grequest(ctx, ds, rq);
}

extern "C" int bnl_set_bge_prior(bnl_ctx *ctx, bnl_data data, double iss_mu, double iss_w,
                                 const double *nu) {
  if (!ctx) return BNL_EINVAL;
  auto it = ctx->data.find(data);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  BnlDataSet &ds = it->second;
  if (!(iss_mu > 0.0)) return bnl_fail(ctx, BNL_EINVAL, "iss.mu must be positive");
  // the prior matrix T = t I needs t > 0, i.e. iss.w > ncol + 1 (check.iss.w,
  // R/sanitization-scores.R:274-291).
  if (iss_w > 0.0 && !(iss_w > (double)ds.ncols + 1.0))
    return bnl_fail(ctx, BNL_EINVAL, "iss.w must be larger than the number of columns + 1");
  ds.bge_mu = iss_mu;
  ds.bge_w = iss_w;
  std::vector<double> by_slot(std::max(ds.nc, 1), 0.0);
  if (nu)
    for (int c = 0; c < ds.ncols; c++) if (ds.is_numeric[c]) by_slot[ds.slot[c]] = nu[c];
  for (size_t d = 0; d < ctx->devs.size(); d++) {
    BnlDevice &dev = ctx->devs[d];
    BnlDeviceData &dd = ds.dev[d];
    BNL_CUDA(ctx, cudaSetDevice(dev.id));
    if (!nu || ds.nc == 0) {
      bnl_pool_free(dev, dd.bge_nu, dd.nu_bytes);
      dd.bge_nu = nullptr; dd.nu_bytes = 0;
      continue;
    }
    if (!dd.bge_nu) {
      dd.nu_bytes = (size_t)ds.nc * sizeof(double);
      int rc = bnl_pool_alloc(ctx, dev, (void **)&dd.bge_nu, dd.nu_bytes);
      if (rc) return rc;
    }
    BNL_CUDA(ctx, cudaMemcpyAsync(dd.bge_nu, by_slot.data(), (size_t)ds.nc * sizeof(double),
                                  cudaMemcpyHostToDevice, dev.stream));
    BNL_CUDA(ctx, cudaStreamSynchronize(dev.stream));
  }
  cudaSetDevice(ctx->devs[0].id);
  return BNL_OK;
}

extern "C" int bnl_set_ebic_gamma(bnl_ctx *ctx, bnl_data data, double gamma) {
  if (!ctx) return BNL_EINVAL;
  auto it = ctx->data.find(data);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  if (!(gamma >= 0.0 && gamma <= 1.0)) return bnl_fail(ctx, BNL_EINVAL, "gamma must lie in [0, 1]");
  it->second.ebic_gamma = gamma;
  return BNL_OK;
}

extern "C" int bnl_gaussian_moments(bnl_ctx *ctx, bnl_data data, double *mean, double *cov) {
  if (!ctx) return BNL_EINVAL;
  auto it = ctx->data.find(data);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  BnlDataSet &ds = it->second;
  if (ds.nc == 0) return bnl_fail(ctx, BNL_EINVAL, "the data set has no numeric columns");
  if (mean) memcpy(mean, ds.h_mean.data(), ds.h_mean.size() * sizeof(double));
  if (cov) memcpy(cov, ds.h_cov.data(), ds.h_cov.size() * sizeof(double));
  return BNL_OK;
}

// Rows sharded over GPUs: every rank registers its block of
// rows, reads its (mean, cov) with bnl_gaussian_moments, the ranks pool them (two
// all-reduces over p and p^2 doubles, bnlearn_b200/replicate.py: pool_moments) and
// each hands the pooled moments back here. From then on every Gaussian test and
// score of this data set is served from them, with nobs_total rows.
extern "C" int bnl_data_set_moments(bnl_ctx *ctx, bnl_data data, int64_t nobs_total, const double *mean,
                                    const double *cov) {
  if (!ctx) return BNL_EINVAL;
  auto it = ctx->data.find(data);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  BnlDataSet &ds = it->second;
  if (ds.nc == 0) return bnl_fail(ctx, BNL_EINVAL, "the data set has no numeric columns");
  if (!mean || !cov || nobs_total < ds.nobs)
    return bnl_fail(ctx, BNL_EINVAL, "bnl_data_set_moments: bad arguments");
  if (ds.any_nan)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "pooled moments need complete numeric data");
  const size_t p = (size_t)ds.nc;
  for (size_t d = 0; d < ds.dev.size(); d++) {
    BNL_CUDA(ctx, cudaSetDevice(ctx->devs[d].id));
    BNL_CUDA(ctx, cudaMemcpy(ds.dev[d].mean, mean, p * sizeof(double), cudaMemcpyHostToDevice));
    BNL_CUDA(ctx, cudaMemcpy(ds.dev[d].cov, cov, p * p * sizeof(double), cudaMemcpyHostToDevice));
  }
  BNL_CUDA(ctx, cudaSetDevice(ctx->devs[0].id));
  ds.h_mean.assign(mean, mean + p);
  ds.h_cov.assign(cov, cov + p * p);
  ds.gauss_nobs = nobs_total;
  return BNL_OK;
}

// ---------------------------------------------------------------------------
// Conditional Gaussian scores (loglik-cg, aic-cg, bic-cg): loglik_cgnode
// (src/scores/cg.loglikelihood.c:119-199) sends every node to the discrete code,
// to the Gaussian code, or -- numeric node with at least one discrete parent --
// to the conditional least squares of ccgloglik (:10-117); the same split here.

namespace {

int run_cg_units(bnl_ctx *ctx, BnlDataSet &ds, std::vector<BnlCgUnit> &units_in, double *score, double *npar,
                 bool nal = false, double kpen = 0.0) {
  const int n = (int)units_in.size();
  if (n == 0) return BNL_OK;
  BnlDevice &dev = ctx->devs[0];
  BnlDeviceData &dd = ds.dev[0];
  BNL_CUDA(ctx, cudaSetDevice(dev.id));
  // Units with a handful of configurations take the register-accumulating kernel,
  // the others the atomic one. A launch handles one kind: the device list holds
  // the first kind, then the second; where[u] is the position of the caller's unit.
  auto small = [](const BnlCgUnit &u) { return u.nconfig <= BNL_CG_SMALL_CFG && u.K <= BNL_CG_SMALL_K; };
  std::vector<BnlCgUnit> units;
  std::vector<int> where(n);
  for (int u = 0; u < n; u++) if (small(units_in[u])) { where[u] = (int)units.size(); units.push_back(units_in[u]); }
  const int nsmall = (int)units.size();
  for (int u = 0; u < n; u++) if (!small(units_in[u])) { where[u] = (int)units.size(); units.push_back(units_in[u]); }
  long long total = 0;
  for (auto &u : units) {
    u.off = total;
    total += (long long)u.nconfig * (1 + u.K + u.K * (u.K + 1) / 2);
  }
  int rc;
  double *mom = nullptr;
  BnlCgUnit *d_units = nullptr;
  double *d_out = nullptr;
  int *d_err = nullptr;
  const size_t mom_bytes = (size_t)total * sizeof(double), unit_bytes = (size_t)n * sizeof(BnlCgUnit);
  const size_t out_bytes = (size_t)n * 4 * sizeof(double);
  if ((rc = bnl_pool_alloc(ctx, dev, (void **)&mom, mom_bytes))) return rc;
  if ((rc = bnl_pool_alloc(ctx, dev, (void **)&d_units, unit_bytes))) { bnl_pool_free(dev, mom, mom_bytes); return rc; }
  if ((rc = bnl_pool_alloc(ctx, dev, (void **)&d_out, out_bytes + 256))) {
    bnl_pool_free(dev, mom, mom_bytes); bnl_pool_free(dev, d_units, unit_bytes);
    return rc;
  }
  d_err = reinterpret_cast<int *>(reinterpret_cast<char *>(d_out) + out_bytes);
  std::vector<double> out((size_t)n * 4);
  int err = 0;
  cudaError_t e = cudaMemsetAsync(mom, 0, mom_bytes, dev.stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_err, 0, sizeof(int), dev.stream);
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(d_units, units.data(), unit_bytes, cudaMemcpyHostToDevice, dev.stream);
  if (e == cudaSuccess) {
    const int chunks = (int)std::max(1ll, std::min(128ll, (ds.nobs + 8191) / 8192));
    const long long chunk_rows = (ds.nobs + chunks - 1) / chunks;
    const size_t smem = BNL_CG_SMEM_DOUBLES * sizeof(double);
    auto pass = [&](int which) {
      if (nsmall > 0) {
        const dim3 gs((unsigned)chunks, (unsigned)nsmall);
        if (which == 1)
          cg_moments_small_kernel<1><<<gs, BNL_CG_THREADS, 0, dev.stream>>>(dd.f64, ds.stride_f64, dd.u8,
              ds.stride_u8, ds.nobs, chunk_rows, d_units, mom, d_err);
        else
          cg_moments_small_kernel<2><<<gs, BNL_CG_THREADS, 0, dev.stream>>>(dd.f64, ds.stride_f64, dd.u8,
              ds.stride_u8, ds.nobs, chunk_rows, d_units, mom, d_err);
      }
      if (n - nsmall > 0) {
        const dim3 ga((unsigned)chunks, (unsigned)(n - nsmall));
        if (which == 1)
          cg_moments_kernel<1><<<ga, BNL_CG_THREADS, smem, dev.stream>>>(dd.f64, ds.stride_f64, dd.u8, ds.stride_u8,
              ds.nobs, chunk_rows, d_units + nsmall, mom, d_err, dd.u16, ds.stride_u16);
        else
          cg_moments_kernel<2><<<ga, BNL_CG_THREADS, smem, dev.stream>>>(dd.f64, ds.stride_f64, dd.u8, ds.stride_u8,
              ds.nobs, chunk_rows, d_units + nsmall, mom, d_err, dd.u16, ds.stride_u16);
      }
    };
    pass(1);
    cg_means_kernel<<<dim3(8, (unsigned)n), 128, 0, dev.stream>>>(d_units, n, mom);
    pass(2);
    cg_score_kernel<<<(n + 63) / 64, 64, 0, dev.stream>>>(d_units, n, mom, d_out, nal ? 1 : 0, kpen, ds.nobs);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(out.data(), d_out, out_bytes, cudaMemcpyDeviceToHost, dev.stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(&err, d_err, sizeof(int), cudaMemcpyDeviceToHost, dev.stream);
  const cudaError_t es = cudaStreamSynchronize(dev.stream);
  bnl_pool_free(dev, mom, mom_bytes);
  bnl_pool_free(dev, d_units, unit_bytes);
  bnl_pool_free(dev, d_out, out_bytes + 256);
  if (e != cudaSuccess || es != cudaSuccess)
    return bnl_fail(ctx, BNL_ECUDA, "conditional Gaussian kernels failed: %s",
                    cudaGetErrorString(e != cudaSuccess ? e : es));
  if (err && !nal)                      // nal-cg / pnal-cg are the scores for incomplete data: rows are skipped
    return bnl_fail(ctx, BNL_EUNSUPPORTED,
                    "conditional Gaussian scores need complete data (a missing value was met)");
  for (int u = 0; u < n; u++) {
    score[u] = out[4 * (size_t)where[u]];
    npar[u] = out[4 * (size_t)where[u] + 1];
  }
  return BNL_OK;
}

}  // namespace

extern "C" int bnl_score_cg(bnl_ctx *ctx, bnl_data data, int score_e, double k, int n, const int *target,
                            const int *pa_off, const int *pa_idx, double *score, double *nparams) {
  if (!ctx) return BNL_EINVAL;
  auto it = ctx->data.find(data);
  if (it == ctx->data.end()) return bnl_fail(ctx, BNL_EINVAL, "unknown data handle");
  BnlDataSet &ds = it->second;
  // conditional Gaussian moments are gathered from the rows this rank holds
  if (ds.gauss_nobs > 0)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "conditional Gaussian scores need the rows, not pooled moments");
  if (score_e == BNL_EBIC_CG) {           // per.node.score.c:194-212
    int rc = bnl_score_cg(ctx, data, BNL_BIC_CG, k, n, target, pa_off, pa_idx, score, nparams);
    if (rc) return rc;
    const double pen = 4.0 * ds.ebic_gamma * log((double)ds.ncols);
    for (int u = 0; u < n; u++) score[u] -= pen * (double)(pa_off[u + 1] - pa_off[u]);
    return BNL_OK;
  }
  const bool nal = score_e == BNL_NAL_CG || score_e == BNL_PNAL_CG;
  if (score_e == BNL_NAL_CG) k = 0.0;     // per.node.score.c:405-408
  if (score_e != BNL_LOGLIK_CG && score_e != BNL_AIC_CG && score_e != BNL_BIC_CG && !nal)
    return bnl_fail(ctx, BNL_EUNSUPPORTED, "score %d is not a conditional Gaussian score handled here", score_e);
  if (n < 0 || (n > 0 && (!target || !pa_off || !score)))
    return bnl_fail(ctx, BNL_EINVAL, "bnl_score_cg: bad arguments");
  // nal_cgnode (src/scores/cg.nal.c:118-200) splits the nodes the same way over
  // nal_dnode_*, glik / cglik_incomplete and ccgnal.
  const int d_code = nal ? BNL_PNAL : score_e == BNL_LOGLIK_CG ? BNL_LOGLIK : (score_e == BNL_AIC_CG ? BNL_AIC : BNL_BIC);
  const int g_code = nal ? BNL_PNAL_G
                         : score_e == BNL_LOGLIK_CG ? BNL_LOGLIK_G : (score_e == BNL_AIC_CG ? BNL_AIC_G : BNL_BIC_G);
  // three sub-requests in the caller's order, plus the nodes that cannot be scored.
  std::vector<int> didx, gidx, cidx;
  std::vector<int> dt, doff(1, 0), dpa, gt, goff(1, 0), gpa;
  std::vector<BnlCgUnit> cg;
  for (int u = 0; u < n; u++) {
    const int t = target[u], p0 = pa_off[u], np = pa_off[u + 1] - p0;
    if (t < 0 || t >= ds.ncols || np < 0)
      return bnl_fail(ctx, BNL_EINVAL, "bnl_score_cg: unit %d refers to a column out of range", u);
    int ndp = 0;
    for (int j = 0; j < np; j++) {
      const int c = pa_idx[p0 + j];
      if (c < 0 || c >= ds.ncols) return bnl_fail(ctx, BNL_EINVAL, "bnl_score_cg: unit %d refers to a column out of range", u);
      ndp += !ds.is_numeric[c];
    }
    if (!ds.is_numeric[t]) {
      if (ndp != np) {           // a discrete node cannot have numeric parents (:155-160)
        score[u] = -INFINITY;
        if (nparams) nparams[u] = 0.0;
        continue;
      }
      didx.push_back(u); dt.push_back(t);
      dpa.insert(dpa.end(), pa_idx + p0, pa_idx + p0 + np);
      doff.push_back((int)dpa.size());
    } else if (ndp == 0) {
      gidx.push_back(u); gt.push_back(t);
      gpa.insert(gpa.end(), pa_idx + p0, pa_idx + p0 + np);
      goff.push_back((int)gpa.size());
    } else {
      BnlCgUnit cu;
      memset(&cu, 0, sizeof(cu));
      if (np - ndp + 1 > BNL_CG_MAXG || ndp > BNL_CG_MAXD)
        return bnl_fail(ctx, BNL_EUNSUPPORTED, "unit %d: at most %d numeric and %d discrete parents", u,
                        BNL_CG_MAXG - 1, BNL_CG_MAXD);
      cu.gslot[cu.K++] = ds.slot[t];
      long long nconfig = 1;
      for (int j = 0; j < np; j++) {
        const int c = pa_idx[p0 + j];
        if (ds.is_numeric[c]) cu.gslot[cu.K++] = ds.slot[c];
        else {
          if (ds.is_wide[c]) cu.dwide |= 1 << cu.nd;          // uint16 block (never the register kernel: >= 128 configurations)
          cu.dslot[cu.nd] = ds.slot[c];
          cu.dmul[cu.nd++] = (int)nconfig;
          nconfig *= ds.nlevels[c];
          if (nconfig >= INT32_MAX / 64)
            return bnl_fail(ctx, BNL_EOVERFLOW, "unit %d: too many configurations of the discrete parents", u);
        }
      }
      cu.nconfig = (int)nconfig;
      cidx.push_back(u);
      cg.push_back(cu);
    }
  }
  int rc;
  std::vector<double> s, p;
  auto scatter = [&](const std::vector<int> &idx) {
    for (size_t i = 0; i < idx.size(); i++) { score[idx[i]] = s[i]; if (nparams) nparams[idx[i]] = p[i]; }
  };
  if (!didx.empty()) {
    s.assign(didx.size(), 0.0); p.assign(didx.size(), 0.0);
    if ((rc = bnl_score_discrete(ctx, data, d_code, k, (int)didx.size(), dt.data(), doff.data(),
                                 dpa.empty() ? nullptr : dpa.data(), s.data(), p.data()))) return rc;
    scatter(didx);
  }
  if (!gidx.empty()) {
    s.assign(gidx.size(), 0.0); p.assign(gidx.size(), 0.0);
    if ((rc = bnl_score_gaussian(ctx, data, g_code, k, (int)gidx.size(), gt.data(), goff.data(),
                                 gpa.empty() ? nullptr : gpa.data(), s.data(), p.data()))) return rc;
    scatter(gidx);
  }
  if (!cidx.empty()) {
    s.assign(cidx.size(), 0.0); p.assign(cidx.size(), 0.0);
    if ((rc = run_cg_units(ctx, ds, cg, s.data(), p.data(), nal, k))) return rc;
    if (score_e != BNL_LOGLIK_CG && !nal)
      for (size_t i = 0; i < cidx.size(); i++) s[i] -= k * p[i];      // per.node.score.c: res - k * nparams
    scatter(cidx);
  }
  return BNL_OK;
}
