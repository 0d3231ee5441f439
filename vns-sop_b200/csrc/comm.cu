// Multi-GPU exchange steps of the path behind the C ABI (north star: "NCCL over NVLink only broadcasts the
// incumbent and reduces the best move"): an NCCL communicator per context, the all-reduce(MIN) of the first-accepted key
// and the broadcast of a path (targets, VNS vector, F / B rows, minima: one contiguous device block) - so a C or C++
// caller needs no torch.  libnccl.so.2 is bound with dlopen at the first vnsgpu_comm_* call: the library itself has no
// link-time dependency on NCCL, and inside a process that already loaded an NCCL (torch) that copy is the one used.
#include <dlfcn.h>

#include "ctx.hpp"

// the few NCCL declarations used (nccl.h 2.x ABI; values checked against /usr/include/nccl.h 2.27)
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef enum { ncclSuccess = 0 } ncclResult_t;
enum { ncclInt8 = 0, ncclChar = 0, ncclInt64 = 4 };
enum { ncclSum = 0, ncclProd = 1, ncclMax = 2, ncclMin = 3 };

struct NcclApi {
   void *lib = nullptr;
   ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
   ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
   ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
   ncclResult_t (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
   ncclResult_t (*Broadcast)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
   ncclResult_t (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
   const char *(*GetErrorString)(ncclResult_t) = nullptr;
   std::string err;
};

static NcclApi &nccl_api()
{
   static NcclApi api;
   if (api.lib || !api.err.empty()) return api;
   const char *names[] = {"libnccl.so.2", "libnccl.so"};
   for (const char *n : names) {
      api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (api.lib) break;
   }
   if (!api.lib) { api.err = std::string("cannot load libnccl.so.2: ") + dlerror(); return api; }
   auto sym = [&](const char *n) { void *p = dlsym(api.lib, n); if (!p && api.err.empty()) api.err = std::string("libnccl lacks ") + n; return p; };
   api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
   api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
   api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
   api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
   api.Broadcast = (decltype(api.Broadcast))sym("ncclBroadcast");
   api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
   api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
   return api;
}

static vnsgpu_status nccl_fail(vnsgpu_ctx *ctx, const char *what, ncclResult_t r)
{
   NcclApi &api = nccl_api();
   ctx->err = std::string(what) + ": " + (api.GetErrorString ? api.GetErrorString(r) : "NCCL error");
   return VNSGPU_E_NCCL;
}
#define NCK(call)                                                    \
   do {                                                              \
      ncclResult_t r__ = (call);                                     \
      if (r__ != ncclSuccess) return nccl_fail(ctx, #call, r__);     \
   } while (0)

// every rank's `bytes` bytes at d_send, in rank order, into d_recv of every rank (best.cu: the ranks' best keys of a sharded
// neighbourhood); a context without communicator copies its own block
vnsgpu_status comm_allgather_bytes(vnsgpu_ctx *ctx, const void *d_send, void *d_recv, size_t bytes, cudaStream_t st)
{
   if (!ctx->comm) {
      CK(cudaMemcpyAsync(d_recv, d_send, bytes, cudaMemcpyDeviceToDevice, st));
      return VNSGPU_OK;
   }
   NCK(nccl_api().AllGather(d_send, d_recv, bytes, ncclChar, (ncclComm_t)ctx->comm, st));
   return VNSGPU_OK;
}

extern "C" {

vnsgpu_status vnsgpu_comm_unique_id(vnsgpu_ctx *ctx, void *id128)
{
   NcclApi &api = nccl_api();
   if (!api.err.empty()) return fail(ctx, VNSGPU_E_NCCL, api.err.c_str());
   ncclUniqueId id;
   NCK(api.GetUniqueId(&id));
   memcpy(id128, &id, sizeof(id));
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_comm_init(vnsgpu_ctx *ctx, const void *id128, int rank, int nranks)
{
   CK(cudaSetDevice(ctx->device));
   if (rank < 0 || rank >= nranks) return fail(ctx, VNSGPU_E_RANGE, "vnsgpu_comm_init: rank outside [0, nranks)");
   if (ctx->comm) return fail(ctx, VNSGPU_E_STATE, "vnsgpu_comm_init: the context already has a communicator");
   NcclApi &api = nccl_api();
   if (!api.err.empty()) return fail(ctx, VNSGPU_E_NCCL, api.err.c_str());
   ncclUniqueId id;
   memcpy(&id, id128, sizeof(id));
   ncclComm_t c = nullptr;
   NCK(api.CommInitRank(&c, nranks, id, rank));
   ctx->comm = c; ctx->comm_owned = true; ctx->comm_rank = rank; ctx->comm_size = nranks;
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_comm_adopt(vnsgpu_ctx *ctx, void *nccl_comm, int rank, int nranks)
{
   if (!nccl_comm || rank < 0 || rank >= nranks) return fail(ctx, VNSGPU_E_RANGE, "vnsgpu_comm_adopt: bad communicator / rank");
   if (ctx->comm) return fail(ctx, VNSGPU_E_STATE, "vnsgpu_comm_adopt: the context already has a communicator");
   NcclApi &api = nccl_api();
   if (!api.err.empty()) return fail(ctx, VNSGPU_E_NCCL, api.err.c_str());
   ctx->comm = nccl_comm; ctx->comm_owned = false; ctx->comm_rank = rank; ctx->comm_size = nranks;
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_comm_destroy(vnsgpu_ctx *ctx)
{
   if (!ctx->comm) return VNSGPU_OK;
   CK(cudaSetDevice(ctx->device));
   CK(cudaStreamSynchronize(ctx->stream));
   if (ctx->comm_owned) NCK(nccl_api().CommDestroy((ncclComm_t)ctx->comm));
   ctx->comm = nullptr; ctx->comm_owned = false; ctx->comm_rank = 0; ctx->comm_size = 1;
   return VNSGPU_OK;
}

vnsgpu_status vnsgpu_comm_info(const vnsgpu_ctx *ctx, int *rank, int *nranks)
{
   if (rank) *rank = ctx->comm_rank;
   if (nranks) *nranks = ctx->comm_size;
   return VNSGPU_OK;
}

// the chosen move: all-reduce(MIN) of the one int64 key the sharded evaluation wrote (index_base + local index, or
// VNSGPU_NO_KEY) - the smallest key is the proposal the reference's sequential loop would have applied
vnsgpu_status vnsgpu_reduce_first_key(vnsgpu_ctx *ctx, long long *d_key, void *stream)
{
   if (!ctx->comm) return ctx->comm_size == 1 ? VNSGPU_OK : fail(ctx, VNSGPU_E_STATE, "no communicator");
   cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
   NCK(nccl_api().AllReduce(d_key, d_key, 1, ncclInt64, ncclMin, (ncclComm_t)ctx->comm, st));
   return VNSGPU_OK;
}

// the incumbent: every rank's `path` becomes a copy of root's (same instance on every rank -> same block layout)
vnsgpu_status vnsgpu_path_broadcast(vnsgpu_ctx *ctx, int path, int root, void *stream)
{
   if (!path_ok(ctx, path)) return VNSGPU_E_STATE;
   if (!ctx->comm) return ctx->comm_size == 1 ? VNSGPU_OK : fail(ctx, VNSGPU_E_STATE, "no communicator");
   if (root < 0 || root >= ctx->comm_size) return fail(ctx, VNSGPU_E_RANGE, "vnsgpu_path_broadcast: root outside the communicator");
   cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
   void *b = ctx->paths[path].block;
   NCK(nccl_api().Broadcast(b, b, ctx->paths[path].bytes, ncclChar, root, (ncclComm_t)ctx->comm, st));
   ctx->paths[path].hdr_ok = false;
   return VNSGPU_OK;
}

// any device buffer (e.g. a generator snapshot the caller keeps on the device)
vnsgpu_status vnsgpu_bcast_bytes(vnsgpu_ctx *ctx, void *d_buf, int64_t nbytes, int root, void *stream)
{
   if (!ctx->comm) return ctx->comm_size == 1 ? VNSGPU_OK : fail(ctx, VNSGPU_E_STATE, "no communicator");
   if (root < 0 || root >= ctx->comm_size || nbytes < 0) return fail(ctx, VNSGPU_E_RANGE, "vnsgpu_bcast_bytes: bad root / size");
   cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
   NCK(nccl_api().Broadcast(d_buf, d_buf, (size_t)nbytes, ncclChar, root, (ncclComm_t)ctx->comm, st));
   return VNSGPU_OK;
}

} // extern "C"
