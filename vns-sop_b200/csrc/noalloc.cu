// vnsgpu_eval_batch_noalloc: the host-buffer candidate evaluation with its scratch (staged descriptors + results) in a
// file-scope __device__ array instead of cudaMalloc'd memory.  No counterpart in the reference; added
// for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc).  The call allocates nothing: no cudaMalloc / cudaFree / cudaMallocHost, no library routine
// that allocates.  It changes no existing code path: it launches its own copy of the block-by-block kernel
// (k_eval_batch, kernels.cuh) and never touches the segment-product / endpoint-vector / sort caches, all of which the
// other entry points allocate on demand.
//
// The array is shared by every call on a device and keeps the last call's contents; the call clears what must start
// clean (the first-accept key) and overwrites the rest before reading it.  Two calls that overlap on the GPU (different
// contexts or streams on one device) race on it: serialise them.
#include "ctx.hpp"
#include "path.cuh"
#include "search.cuh"
#include "eval.cuh"
#include "kernels.cuh"

// per candidate: 16 B descriptor + 8 B length + 8 B lower bound + 1 B flags; + 256 B header (first-accept key)
#define NA_HDR 256
#define NA_BYTES (NA_HDR + (size_t)VNSGPU_NOALLOC_MAX_CANDS * (16 + 8 + 8) + (((size_t)VNSGPU_NOALLOC_MAX_CANDS + 255) & ~(size_t)255))

static __device__ __align__(256) unsigned char g_na_scratch[NA_BYTES];   // ~0.5 MB on every GPU that loads the library, until exit

// In host code `g_na_scratch` is only a stand-in for the symbol, not a device address: the address a kernel may use
// comes from cudaGetSymbolAddress, once per device, with that device current.
#define NA_MAX_DEVICES 64
static unsigned char *g_na_ptr[NA_MAX_DEVICES];

static vnsgpu_status na_scratch(vnsgpu_ctx *ctx, unsigned char **out)
{
   const int dev = ctx->device;
   if (dev < 0 || dev >= NA_MAX_DEVICES) return fail(ctx, VNSGPU_E_RANGE, "vnsgpu_eval_batch_noalloc: device ordinal beyond 63");
   if (!g_na_ptr[dev]) {
      void *p = nullptr;
      CK(cudaGetSymbolAddress(&p, g_na_scratch));     // the current device is ctx->device (set by the caller below)
      g_na_ptr[dev] = (unsigned char *)p;
   }
   *out = g_na_ptr[dev];
   return VNSGPU_OK;
}

extern "C" {

int64_t vnsgpu_noalloc_max_cands(void) { return VNSGPU_NOALLOC_MAX_CANDS; }

vnsgpu_status vnsgpu_eval_batch_noalloc(vnsgpu_ctx *ctx, int path, int64_t n, const vnsgpu_cand *cands, double budget, int skip,
                                        double *len, double *lb, uint8_t *flags, int64_t *first_accept)
{
   CK(cudaSetDevice(ctx->device));
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (first_accept) *first_accept = -1;
   if (n <= 0) return VNSGPU_OK;
   if (n > VNSGPU_NOALLOC_MAX_CANDS) {
      ctx->err = "vnsgpu_eval_batch_noalloc: n = " + std::to_string((long long)n) + " exceeds VNSGPU_NOALLOC_MAX_CANDS = " +
                 std::to_string((long long)VNSGPU_NOALLOC_MAX_CANDS) + " (fixed static scratch; split the batch or use vnsgpu_eval_batch)";
      return VNSGPU_E_RANGE;
   }
   const size_t smem = sizeof(double) * 2 * ctx->I.mp * (VNS_EVAL_THREADS / 32);
   if (smem > 48 * 1024) return fail(ctx, VNSGPU_E_TOO_LARGE, "vnsgpu_eval_batch_noalloc: clusters beyond 384 samples");
   unsigned char *b = nullptr;
   vnsgpu_status s = na_scratch(ctx, &b);
   if (s) return s;
   unsigned long long *d_first = (unsigned long long *)b;
   vnsgpu_cand *d_c = (vnsgpu_cand *)(b + NA_HDR);
   double *d_len = (double *)(b + NA_HDR + (size_t)VNSGPU_NOALLOC_MAX_CANDS * 16);
   double *d_lb = d_len + VNSGPU_NOALLOC_MAX_CANDS;
   uint8_t *d_fl = (uint8_t *)(d_lb + VNSGPU_NOALLOC_MAX_CANDS);
   cudaStream_t st = ctx->stream;
   CK(cudaMemsetAsync(d_first, 0xFF, sizeof(unsigned long long), st));          // -1 = none; everything else is overwritten
   CK(cudaMemcpyAsync(d_c, cands, sizeof(vnsgpu_cand) * (size_t)n, cudaMemcpyHostToDevice, st));
   const int64_t wpb = VNS_EVAL_THREADS / 32;
   const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n + wpb - 1) / wpb, (int64_t)ctx->num_sms * 8));
   k_eval_batch<<<grid, VNS_EVAL_THREADS, smem, st>>>(ctx->I, ctx->paths[path].d, n, 0, d_c, budget, skip, d_len, d_lb, d_fl, d_first);
   LAUNCHED(ctx);
   if (len) CK(cudaMemcpyAsync(len, d_len, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, st));
   if (lb) CK(cudaMemcpyAsync(lb, d_lb, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, st));
   if (flags) CK(cudaMemcpyAsync(flags, d_fl, (size_t)n, cudaMemcpyDeviceToHost, st));
   long long first = -1;
   if (first_accept) CK(cudaMemcpyAsync(&first, d_first, sizeof(long long), cudaMemcpyDeviceToHost, st));
   CK(cudaStreamSynchronize(st));
   if (first_accept) *first_accept = (int64_t)first;
   return VNSGPU_OK;
}

} // extern "C"
