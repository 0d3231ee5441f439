// Measurement support (bench.py's roofline block): streaming-read micro-benchmarks for the level that holds a kernel's
// working set - L1 (every thread block re-reads its own 64 KB), L2 (all blocks re-read one 32 MB buffer), HBM (a buffer
// far beyond L2 read once per pass) - and the per-phase device times of the batched evaluation call (CUDA events on the
// launching stream; the hooks are in api_path.cu).  No reference counterpart: SURVEY.md 8(d) asks the builder to measure
// the roof that binds instead of quoting a nominal one.
#include "ctx.hpp"

// every thread sums 16-byte loads; `span` bytes per block (L1 form) or for the whole grid (L2 / HBM form)
template <bool PER_BLOCK>
static __global__ void __launch_bounds__(512) k_stream_read(const int4 *__restrict__ buf, size_t span_vec, int passes, int *sink)
{
   int4 acc = make_int4(0, 0, 0, 0);
   if (PER_BLOCK) {
      const int4 *mine = buf + (size_t)blockIdx.x * span_vec;
      for (int p = 0; p < passes; ++p)
         for (size_t i = threadIdx.x; i < span_vec; i += blockDim.x) {
            int4 v;
            asm volatile("ld.global.ca.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mine + i));
            acc.x += v.x; acc.y ^= v.y; acc.z += v.z; acc.w ^= v.w;
         }
   } else {
      const size_t stride = (size_t)gridDim.x * blockDim.x;
      for (int p = 0; p < passes; ++p)
         for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < span_vec; i += stride) {
            int4 v;
            asm volatile("ld.global.cg.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(buf + i));
            acc.x += v.x; acc.y ^= v.y; acc.z += v.z; acc.w ^= v.w;
         }
   }
   if ((acc.x ^ acc.y ^ acc.z ^ acc.w) == 0x7fffffff) *sink = 1;   // never true for the zeroed buffer: keeps the loads alive
}

extern "C" {

// gbs[0] = L1-resident, gbs[1] = L2-resident, gbs[2] = HBM streaming read, GB/s over the whole device (best of 5).
// Allocates and frees its own buffers (4 GiB for the HBM pass; VNSGPU_E_NOMEM when that does not fit).
vnsgpu_status vnsgpu_measure_read_peaks(vnsgpu_ctx *ctx, double *gbs)
{
   CK(cudaSetDevice(ctx->device));
   const int nb = ctx->num_sms * 2;
   const size_t l1_span = 64 << 10, l2_bytes = (size_t)32 << 20, hbm_bytes = (size_t)4 << 30;
   char *buf = nullptr;
   int *sink = nullptr;
   if (cudaMalloc(&buf, hbm_bytes) != cudaSuccess) { cudaGetLastError(); return fail(ctx, VNSGPU_E_NOMEM, "vnsgpu_measure_read_peaks: 4 GiB buffer"); }
   cudaError_t e = cudaMalloc(&sink, sizeof(int));
   if (e == cudaSuccess) e = cudaMemsetAsync(buf, 0, hbm_bytes, ctx->stream);
   cudaEvent_t e0 = nullptr, e1 = nullptr;
   if (e == cudaSuccess) e = cudaEventCreate(&e0);
   if (e == cudaSuccess) e = cudaEventCreate(&e1);
   for (int which = 0; which < 3 && e == cudaSuccess; ++which) {
      double best = 0;
      for (int rep = 0; rep < 6 && e == cudaSuccess; ++rep) {      // rep 0 warms the cache level
         const int passes = which == 0 ? 400 : (which == 1 ? 40 : 1);
         cudaEventRecord(e0, ctx->stream);
         if (which == 0) k_stream_read<true><<<nb, 512, 0, ctx->stream>>>((const int4 *)buf, l1_span / 16, passes, sink);
         else k_stream_read<false><<<nb * 2, 512, 0, ctx->stream>>>((const int4 *)buf, (which == 1 ? l2_bytes : hbm_bytes) / 16, passes, sink);
         ctx->launches++;
         cudaEventRecord(e1, ctx->stream);
         e = cudaStreamSynchronize(ctx->stream);
         float ms = 0;
         if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
         const double bytes = which == 0 ? (double)l1_span * nb * passes : (which == 1 ? (double)l2_bytes * passes : (double)hbm_bytes);
         if (rep > 0 && ms > 0) best = std::max(best, bytes / (ms * 1e-3) / 1e9);
      }
      gbs[which] = best;
   }
   if (e0) cudaEventDestroy(e0);
   if (e1) cudaEventDestroy(e1);
   cudaFree(sink);
   cudaFree(buf);
   if (e != cudaSuccess) { ctx->err = std::string("vnsgpu_measure_read_peaks: ") + cudaGetErrorString(e); return VNSGPU_E_CUDA; }
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_set_phase_timing(vnsgpu_ctx *ctx, int enable)
{
   CK(cudaSetDevice(ctx->device));
   CK(cudaStreamSynchronize(ctx->stream));
   ctx->phase_timing = enable != 0;
   for (auto &pe : ctx->phase_events) { cudaEventDestroy(pe.e0); cudaEventDestroy(pe.e1); }
   ctx->phase_events.clear();
   return VNSGPU_OK;
}

// ms[k] / count[k]: device time and number of timed intervals of phase k since the last call (VNSGPU_PHASE_*).  The
// caller synchronises the stream(s) the evaluation calls ran on first.
vnsgpu_status vnsgpu_phase_times(vnsgpu_ctx *ctx, double *ms, int64_t *count)
{
   CK(cudaSetDevice(ctx->device));
   for (int k = 0; k < VNSGPU_PHASES; ++k) { ms[k] = 0; count[k] = 0; }
   for (auto &pe : ctx->phase_events) {
      float t = 0;
      CK(cudaEventSynchronize(pe.e1));
      CK(cudaEventElapsedTime(&t, pe.e0, pe.e1));
      ms[pe.phase] += t; count[pe.phase] += 1;
      cudaEventDestroy(pe.e0); cudaEventDestroy(pe.e1);
   }
   ctx->phase_events.clear();
   return VNSGPU_OK;
}

} // extern "C"
