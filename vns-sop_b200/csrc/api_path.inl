// C ABI: path handles + batched evaluation (included by vnsgpu.cu)

static bool path_ok(vnsgpu_ctx *ctx, int path)
{
   if (!ctx->have_table) { ctx->err = "no distance table"; return false; }
   if (path < 0 || path >= (int)ctx->paths.size()) { ctx->err = "bad path handle"; return false; }
   return true;
}

static size_t al16(size_t x) { return (x + 15) & ~(size_t)15; }

// host mirror of the path header (read back after every mutating call)
struct PathHdr {
   int hdr[HDR_INTS];
   double hd[HD_DBLS];
};

static vnsgpu_status read_hdr(vnsgpu_ctx *ctx, int path, PathHdr *h)
{
   const PathDev &d = ctx->paths[path].d;
   // hdr and hd are adjacent at the start of the block
   CK(cudaMemcpyAsync(h, d.hdr, sizeof(PathHdr), cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

static vnsgpu_status check_path_err(vnsgpu_ctx *ctx, int path)
{
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   if (h.hdr[HDR_ERR] == 4) return fail(ctx, VNSGPU_E_INCONSISTENT, "forward/backward path length mismatch");
   if (h.hdr[HDR_ERR] == 3) return fail(ctx, VNSGPU_E_STATE, "proposal rejection loop did not terminate");
   if (h.hdr[HDR_ERR]) return fail(ctx, VNSGPU_E_STATE, "device-side error flag");
   return VNSGPU_OK;
}

static size_t chain_smem(const vnsgpu_ctx *ctx) { return sizeof(double) * 2 * ctx->I.mp * VNS_CHAIN_WARPS; }

extern "C" {

vnsgpu_status vnsgpu_path_create(vnsgpu_ctx *ctx, int *path)
{
   CK(cudaSetDevice(ctx->device));
   if (!ctx->have_table) return fail(ctx, VNSGPU_E_STATE, "build or set the distance table first");
   const InstDev &I = ctx->I;
   const size_t rows = I.rows, mp = I.mp;
   size_t o_hdr = 0;
   size_t o_hd = o_hdr + sizeof(int) * HDR_INTS;
   size_t o_tg = al16(o_hd + sizeof(double) * HD_DBLS);
   size_t o_vns = al16(o_tg + sizeof(int) * rows);
   size_t o_F = al16(o_vns + sizeof(int) * rows);
   size_t o_B = al16(o_F + sizeof(double) * rows * mp);
   size_t o_minF = al16(o_B + sizeof(double) * rows * mp);
   size_t o_minB = al16(o_minF + sizeof(double) * rows);
   size_t total = al16(o_minB + sizeof(double) * rows);
   PathHost ph;
   CK(cudaMalloc(&ph.block, total));
   CK(cudaMemsetAsync(ph.block, 0, total, ctx->stream));
   char *b = (char *)ph.block;
   ph.d.hdr = (int *)(b + o_hdr); ph.d.hd = (double *)(b + o_hd);
   ph.d.tg = (int *)(b + o_tg); ph.d.vns = (int *)(b + o_vns);
   ph.d.F = (double *)(b + o_F); ph.d.B = (double *)(b + o_B);
   ph.d.minF = (double *)(b + o_minF); ph.d.minB = (double *)(b + o_minB);
   ctx->paths.push_back(ph);
   *path = (int)ctx->paths.size() - 1;
   k_path_update<<<1, 64, 0, ctx->stream>>>(ctx->I, ph.d);   // empty path start -> goal
   LAUNCHED(ctx);
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_path_copy(vnsgpu_ctx *ctx, int dst, int src)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, dst) || !path_ok(ctx, src)) return VNSGPU_E_STATE;
   k_path_copy<<<1, 256, 0, ctx->stream>>>(ctx->I, ctx->paths[dst].d, ctx->paths[src].d);
   LAUNCHED(ctx);
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_path_set(vnsgpu_ctx *ctx, int path, const int *vns, int nvns, int T)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   const InstDev &I = ctx->I;
   if (nvns < 0 || nvns > I.C || T < 0 || T > nvns) return fail(ctx, VNSGPU_E_RANGE, "bad T / nvns");
   for (int k = 0; k < nvns; ++k)
      if (vns[k] < 0 || vns[k] >= I.C) return fail(ctx, VNSGPU_E_RANGE, "cluster id out of range");
   const PathDev &d = ctx->paths[path].d;
   int hdr[HDR_INTS] = {T, nvns, 0, 0, 0, 0, 0, 0};
   CK(cudaMemcpyAsync(d.hdr, hdr, sizeof(hdr), cudaMemcpyHostToDevice, ctx->stream));
   if (nvns) CK(cudaMemcpyAsync(d.vns, vns, sizeof(int) * nvns, cudaMemcpyHostToDevice, ctx->stream));
   if (T) CK(cudaMemcpyAsync(d.tg, vns, sizeof(int) * T, cudaMemcpyHostToDevice, ctx->stream));
   k_path_update<<<1, 64, 0, ctx->stream>>>(ctx->I, d);
   LAUNCHED(ctx);
   return check_path_err(ctx, path);
}

vnsgpu_status vnsgpu_path_add_point(vnsgpu_ctx *ctx, int path, int cluster, int idx)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   const int T = h.hdr[HDR_T];
   if (idx > T || idx < 0) return fail(ctx, VNSGPU_E_RANGE, "addPoint idx exceeds maximum");   // vns_sop_path.cpp:69-72
   if (cluster < 0 || cluster >= ctx->I.C || T + 3 > ctx->I.rows) return fail(ctx, VNSGPU_E_RANGE, "addPoint cluster / capacity");
   k_path_edit<<<1, 256, 0, ctx->stream>>>(ctx->I, ctx->paths[path].d, 0, cluster, idx);
   LAUNCHED(ctx);
   return check_path_err(ctx, path);
}

vnsgpu_status vnsgpu_path_remove_point(vnsgpu_ctx *ctx, int path, int idx)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   if (idx >= h.hdr[HDR_T] || idx < 0) return fail(ctx, VNSGPU_E_RANGE, "removePoint idx exceeds maximum");   // :158-161
   k_path_edit<<<1, 256, 0, ctx->stream>>>(ctx->I, ctx->paths[path].d, 1, 0, idx);
   LAUNCHED(ctx);
   return check_path_err(ctx, path);
}

vnsgpu_status vnsgpu_path_get(vnsgpu_ctx *ctx, int path, int *T, int *targets, int *nvns, int *vns, double *length,
                              double *length_rear, double *reward)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   const PathDev &d = ctx->paths[path].d;
   if (T) *T = h.hdr[HDR_T];
   if (nvns) *nvns = h.hdr[HDR_NVNS];
   if (length) *length = h.hd[HD_LEN];
   if (length_rear) *length_rear = h.hd[HD_LEN_REAR];
   if (reward) *reward = h.hd[HD_REWARD];
   if (targets && h.hdr[HDR_T]) CK(cudaMemcpyAsync(targets, d.tg, sizeof(int) * h.hdr[HDR_T], cudaMemcpyDeviceToHost, ctx->stream));
   if (vns && h.hdr[HDR_NVNS]) CK(cudaMemcpyAsync(vns, d.vns, sizeof(int) * h.hdr[HDR_NVNS], cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_path_get_solution(vnsgpu_ctx *ctx, int path, int *cluster, int *node, double *length_calc)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   const int n = h.hdr[HDR_T] + 2;
   s = need_scratch(ctx, sizeof(int) * 2 * n + 16);
   if (s) return s;
   double *d_calc = (double *)ctx->d_scratch;
   int *d_cl = (int *)((char *)ctx->d_scratch + 16), *d_nd = d_cl + n;
   k_path_solution<<<1, 32, 0, ctx->stream>>>(ctx->I, ctx->paths[path].d, d_cl, d_nd, d_calc);
   LAUNCHED(ctx);
   if (cluster) CK(cudaMemcpyAsync(cluster, d_cl, sizeof(int) * n, cudaMemcpyDeviceToHost, ctx->stream));
   if (node) CK(cudaMemcpyAsync(node, d_nd, sizeof(int) * n, cudaMemcpyDeviceToHost, ctx->stream));
   if (length_calc) CK(cudaMemcpyAsync(length_calc, d_calc, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_path_get_dp(vnsgpu_ctx *ctx, int path, double *F, double *B, int stride)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (stride < ctx->I.mp) return fail(ctx, VNSGPU_E_RANGE, "stride < max cluster size");
   PathHdr h;
   vnsgpu_status s = read_hdr(ctx, path, &h);
   if (s) return s;
   const PathDev &d = ctx->paths[path].d;
   const size_t rows = h.hdr[HDR_T] + 2, w = sizeof(double) * ctx->I.mp;
   if (F) CK(cudaMemcpy2DAsync(F, sizeof(double) * stride, d.F, w, w, rows, cudaMemcpyDeviceToHost, ctx->stream));
   if (B) CK(cudaMemcpy2DAsync(B, sizeof(double) * stride, d.B, w, w, rows, cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   return VNSGPU_OK;
}

static int eval_grid(const vnsgpu_ctx *ctx, int64_t n)
{
   const int64_t wpb = VNS_EVAL_THREADS / 32;
   const int64_t need = (n + wpb - 1) / wpb;
   const int64_t cap = (int64_t)ctx->num_sms * 8;   // 8 resident CTAs of 256 threads per SM
   return (int)std::max<int64_t>(1, std::min(need, cap));
}

vnsgpu_status vnsgpu_eval_batch_dev(vnsgpu_ctx *ctx, int path, int64_t n, const vnsgpu_cand *d_cands, double budget, int skip,
                                    double *d_len, double *d_lb, uint8_t *d_flags, long long *d_first_accept, void *stream)
{
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (n <= 0) return VNSGPU_OK;
   cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
   const size_t smem = sizeof(double) * 2 * ctx->I.mp * (VNS_EVAL_THREADS / 32);
   if (d_first_accept) CK(cudaMemsetAsync(d_first_accept, 0xFF, sizeof(long long), st));   // -1 = none
   k_eval_batch<<<eval_grid(ctx, n), VNS_EVAL_THREADS, smem, st>>>(ctx->I, ctx->paths[path].d, n, d_cands, budget, skip, d_len, d_lb,
                                                                    d_flags, (unsigned long long *)d_first_accept);
   LAUNCHED(ctx);
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_eval_batch(vnsgpu_ctx *ctx, int path, int64_t n, const vnsgpu_cand *cands, double budget, int skip, double *len,
                                double *lb, uint8_t *flags, int64_t *first_accept)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (first_accept) *first_accept = -1;
   if (n <= 0) return VNSGPU_OK;
   const size_t o_len = al16(sizeof(vnsgpu_cand) * n), o_lb = o_len + al16(sizeof(double) * n);
   const size_t o_fl = o_lb + al16(sizeof(double) * n);
   vnsgpu_status s = need_scratch(ctx, o_fl + al16(n));
   if (s) return s;
   char *b = (char *)ctx->d_scratch;
   vnsgpu_cand *d_c = (vnsgpu_cand *)b;
   double *d_len = (double *)(b + o_len), *d_lb = (double *)(b + o_lb);
   uint8_t *d_fl = (uint8_t *)(b + o_fl);
   CK(cudaMemcpyAsync(d_c, cands, sizeof(vnsgpu_cand) * n, cudaMemcpyHostToDevice, ctx->stream));
   s = vnsgpu_eval_batch_dev(ctx, path, n, d_c, budget, skip, len ? d_len : nullptr, lb ? d_lb : nullptr, flags ? d_fl : nullptr,
                             ctx->d_first, ctx->stream);
   if (s) return s;
   long long first = 0;
   if (len) CK(cudaMemcpyAsync(len, d_len, sizeof(double) * n, cudaMemcpyDeviceToHost, ctx->stream));
   if (lb) CK(cudaMemcpyAsync(lb, d_lb, sizeof(double) * n, cudaMemcpyDeviceToHost, ctx->stream));
   if (flags) CK(cudaMemcpyAsync(flags, d_fl, n, cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaMemcpyAsync(&first, ctx->d_first, sizeof(first), cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   if (first_accept) *first_accept = (int64_t)first;
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_cand_bytes(vnsgpu_ctx *ctx, int path, int64_t n, const vnsgpu_cand *cands, int64_t *bytes)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   *bytes = 0;
   if (n <= 0) return VNSGPU_OK;
   vnsgpu_status s = need_scratch(ctx, sizeof(vnsgpu_cand) * n + 16);
   if (s) return s;
   unsigned long long *d_tot = (unsigned long long *)ctx->d_scratch;
   vnsgpu_cand *d_c = (vnsgpu_cand *)((char *)ctx->d_scratch + 16);
   CK(cudaMemsetAsync(d_tot, 0, 8, ctx->stream));
   CK(cudaMemcpyAsync(d_c, cands, sizeof(vnsgpu_cand) * n, cudaMemcpyHostToDevice, ctx->stream));
   const int grid = (int)std::min<int64_t>((n + 255) / 256, 1184);
   k_cand_bytes<<<grid, 256, 0, ctx->stream>>>(ctx->I, ctx->paths[path].d, n, d_c, d_tot);
   LAUNCHED(ctx);
   unsigned long long tot = 0;
   CK(cudaMemcpyAsync(&tot, d_tot, 8, cudaMemcpyDeviceToHost, ctx->stream));
   CK(cudaStreamSynchronize(ctx->stream));
   *bytes = (int64_t)tot;
   return VNSGPU_OK;
}

} // extern "C"
