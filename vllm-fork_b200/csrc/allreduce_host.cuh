// b200ops -- host side of the NVLink all-reduce (included by allreduce.cu): IPC helpers, context,
// launchers and the C ABI. Mirrors the life cycle of the reference's CustomAllreduce
// (custom_all_reduce.py:70-260 init / create_shared_buffer / register_buffer / close;
// csrc/custom_all_reduce.cu init_custom_ar / register_buffer / all_reduce / dispose).
#pragma once

namespace b200 {
namespace ar {

static int find_region(const Context* c, const void* ptr, size_t bytes, size_t* off) {
  const uint8_t* p = reinterpret_cast<const uint8_t*>(ptr);
  for (int i = 0; i < c->num_regions; i++) {
    const uint8_t* b = c->regions[i].base[c->rank];
    if (p >= b && p + bytes <= b + c->regions[i].bytes) {
      *off = (size_t)(p - b);
      return i;
    }
  }
  return -1;
}

// Fill `pe` for an input of `bytes` bytes at `in`: zero-copy if it lives in a registered region,
// else staged (returns the pointer the kernel must copy from in *stage_src).
static int resolve_input(const Context* c, const void* in, size_t bytes, Peers* pe, const uint8_t** stage_src,
                         const char* op) {
  for (int r = 0; r < c->world; r++) {
    pe->sig[r] = reinterpret_cast<Signal*>(c->regions[0].base[r]);
    pe->tmp[r] = c->regions[0].base[r] + c->staging_off + c->staging_bytes;  // scratch follows staging
  }
  size_t off;
  pe->mc_in = nullptr;
  pe->pdl_mode = pdl_collective_mode();
  pe->spin_ns = spin_timeout_ns();
  const int reg = find_region(c, in, bytes, &off);
  if (reg >= 0) {
    // region 0: the user heap starts after signals + staging + scratch; staging/scratch themselves are
    // not valid inputs (the kernels overwrite them)
    if (reg == 0 && off < c->staging_off + 2 * c->staging_bytes) {
      set_last_error("%s: input overlaps the all-reduce staging area", op);
      return B200_ERR_INVALID;
    }
    for (int r = 0; r < c->world; r++) pe->in[r] = c->regions[reg].base[r] + off;
    if (reg == 0 && c->mc_base) pe->mc_in = c->mc_base + off;
    *stage_src = nullptr;
  } else {
    if (bytes > c->staging_bytes) {
      set_last_error("%s: unregistered input of %zu bytes exceeds the staging area (%zu bytes)", op, bytes,
                     c->staging_bytes);
      return B200_ERR_WORKSPACE;
    }
    for (int r = 0; r < c->world; r++) pe->in[r] = c->regions[0].base[r] + c->staging_off;
    if (c->mc_base) pe->mc_in = c->mc_base + c->staging_off;
    *stage_src = reinterpret_cast<const uint8_t*>(in);
  }
  return B200_OK;
}

// collectives: programmatic launch unless B200_PDL_AR=2 (then a plain launch: the kernel starts after its predecessor
// has completed and lets its successor start only when it has completed itself)
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_ar_kernel(void (*kern)(KArgs...), dim3 grid, dim3 block, cudaStream_t stream,
                                           Args&&... args) {
  if (pdl_collective_mode() == 2) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.stream = stream;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
  }
  return launch_pdl(kern, grid, block, 0, stream, std::forward<Args>(args)...);
}

template <typename T, int N>
static int launch_ar(const Peers& pe, void* out, const uint8_t* stage_src, int rank, int64_t packs, bool two_shot,
                     int skip_end, cudaStream_t stream) {
  // one trip of the unrolled loop per thread where possible; never more CTAs than SMs (1 CTA / SM,
  // all CTAs must be co-resident: they meet in the barriers)
  const int per_cta = NTHREADS * (two_shot ? 2 * N : (N <= 4 ? 4 : 2));
  int64_t blocks = (packs + per_cta - 1) / per_cta;
  const int cap = sm_count() < MAX_BLOCKS ? sm_count() : MAX_BLOCKS;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (two_shot)
    B200_CUDA(launch_ar_kernel(allreduce_two_shot_kernel<T, N>, dim3((unsigned)blocks), dim3(NTHREADS), stream, pe,
                               (uint8_t*)out, stage_src, rank, packs, skip_end));
  else
    B200_CUDA(launch_ar_kernel(allreduce_one_shot_kernel<T, N>, dim3((unsigned)blocks), dim3(NTHREADS), stream, pe,
                               (uint8_t*)out, stage_src, rank, packs, skip_end));
  return B200_OK;
}

template <typename T>
static int dispatch_ar(int world, const Peers& pe, void* out, const uint8_t* stage_src, int rank, int64_t packs,
                       bool two_shot, int skip_end, cudaStream_t stream) {
  switch (world) {
    case 2: return launch_ar<T, 2>(pe, out, stage_src, rank, packs, two_shot, skip_end, stream);
    case 4: return launch_ar<T, 4>(pe, out, stage_src, rank, packs, two_shot, skip_end, stream);
    case 6: return launch_ar<T, 6>(pe, out, stage_src, rank, packs, two_shot, skip_end, stream);
    case 8: return launch_ar<T, 8>(pe, out, stage_src, rank, packs, two_shot, skip_end, stream);
    default:
      set_last_error("custom allreduce only supports num gpus in (2,4,6,8), got %d", world);
      return B200_ERR_UNSUPPORTED;
  }
}

template <typename T, int N>
static int launch_mc(const Peers& pe, void* out, uint8_t* mc_out, const uint8_t* stage_src, int rank, int64_t packs,
                     cudaStream_t stream) {
  const int64_t mine = mc_out ? (packs + N - 1) / N : packs;
  int64_t blocks = (mine + NTHREADS - 1) / NTHREADS;    // one 16-byte in-switch reduction per thread where possible
  const int cap = sm_count() < MAX_BLOCKS ? sm_count() : MAX_BLOCKS;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  B200_CUDA(launch_ar_kernel(allreduce_multimem_kernel<T, N>, dim3((unsigned)blocks), dim3(NTHREADS), stream, pe,
                             (uint8_t*)out, mc_out, stage_src, rank, packs));
  return B200_OK;
}

template <typename T>
static int dispatch_mc(int world, const Peers& pe, void* out, uint8_t* mc_out, const uint8_t* stage_src, int rank,
                       int64_t packs, cudaStream_t stream) {
  switch (world) {
    case 2: return launch_mc<T, 2>(pe, out, mc_out, stage_src, rank, packs, stream);
    case 4: return launch_mc<T, 4>(pe, out, mc_out, stage_src, rank, packs, stream);
    case 6: return launch_mc<T, 6>(pe, out, mc_out, stage_src, rank, packs, stream);
    case 8: return launch_mc<T, 8>(pe, out, mc_out, stage_src, rank, packs, stream);
    default:
      set_last_error("custom allreduce only supports num gpus in (2,4,6,8), got %d", world);
      return B200_ERR_UNSUPPORTED;
  }
}

template <typename T, int N, bool QUANT>
static int launch_fused(const Peers& pe, const FusedParams& fp, const uint8_t* stage_src, int rank,
                        cudaStream_t stream) {
  const int nv = fp.hidden / (16 / (int)sizeof(T));
  int blocks = fp.rows_per_rank < 1 ? 1 : fp.rows_per_rank;
  if (blocks > MAX_BLOCKS) blocks = MAX_BLOCKS;
#define B200_FUSED(RR)                                                                                   \
  do {                                                                                                   \
    if (fp.mc_out && pe.mc_in)                                                                           \
      B200_CUDA(launch_ar_kernel(allreduce_fused_norm_kernel<T, N, QUANT, RR, true>, dim3((unsigned)blocks),    \
                                 dim3(256), stream, pe, fp, stage_src, rank));                            \
    else                                                                                                 \
      B200_CUDA(launch_ar_kernel(allreduce_fused_norm_kernel<T, N, QUANT, RR, false>, dim3((unsigned)blocks),   \
                                 dim3(256), stream, pe, fp, stage_src, rank));                            \
  } while (0)
  if (nv <= 256) B200_FUSED(1);
  else if (nv <= 512) B200_FUSED(2);
  else if (nv <= 1024) B200_FUSED(4);
  else B200_FUSED(8);
#undef B200_FUSED
  B200_LAUNCH_CHECK();
  return B200_OK;
}

template <typename T, bool QUANT>
static int dispatch_fused(int world, const Peers& pe, const FusedParams& fp, const uint8_t* stage_src, int rank,
                          cudaStream_t stream) {
  switch (world) {
    case 2: return launch_fused<T, 2, QUANT>(pe, fp, stage_src, rank, stream);
    case 4: return launch_fused<T, 4, QUANT>(pe, fp, stage_src, rank, stream);
    case 8: return launch_fused<T, 8, QUANT>(pe, fp, stage_src, rank, stream);
    default:
      set_last_error("fused all-reduce supports 2, 4 or 8 ranks, got %d", world);
      return B200_ERR_UNSUPPORTED;
  }
}

}  // namespace ar
}  // namespace b200

using namespace b200;
using namespace b200::ar;

// ---------------------------------------------------------------- IPC helpers
extern "C" int b200_ipc_alloc(size_t bytes, void** ptr, void* handle64) {
  B200_CHECK(ptr && handle64 && bytes > 0, B200_ERR_INVALID, "b200_ipc_alloc: bad arguments");
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "b200_ipc_alloc: no CUDA device (no CPU fallback)");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  B200_CUDA(cudaMalloc(ptr, bytes));
  B200_CUDA(cudaMemset(*ptr, 0, bytes));
  B200_CUDA(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  B200_CUDA(cudaIpcGetMemHandle(&h, *ptr));
  memcpy(handle64, &h, 64);
  return B200_OK;
}
extern "C" int b200_ipc_open(const void* handle64, void** ptr) {
  B200_CHECK(ptr && handle64, B200_ERR_INVALID, "b200_ipc_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  B200_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return B200_OK;
}
extern "C" int b200_ipc_close(void* ptr) {
  B200_CUDA(cudaIpcCloseMemHandle(ptr));
  return B200_OK;
}
extern "C" int b200_ipc_free(void* ptr) {
  B200_CUDA(cudaFree(ptr));
  return B200_OK;
}

// ---------------------------------------------------------------- context
extern "C" size_t b200_allreduce_meta_bytes(void) { return SIGNAL_BYTES; }

extern "C" int b200_allreduce_init(void** ctx_out, int rank, int world, void* const* buffer_ptrs,
                                   size_t buffer_bytes, size_t max_message_bytes) {
  B200_CHECK(ctx_out && buffer_ptrs, B200_ERR_INVALID, "allreduce_init: null pointer");
  B200_CHECK(world >= 2 && world <= MAX_RANKS && world % 2 == 0, B200_ERR_UNSUPPORTED,
             "custom allreduce only supports num gpus in (2,4,6,8), got %d", world);
  B200_CHECK(rank >= 0 && rank < world, B200_ERR_INVALID, "allreduce_init: invalid rank passed in");
  B200_CHECK(max_message_bytes % 1024 == 0 && max_message_bytes > 0, B200_ERR_INVALID,
             "allreduce_init: max_message_bytes must be a positive multiple of 1024");
  B200_CHECK(buffer_bytes >= SIGNAL_BYTES + 2 * max_message_bytes, B200_ERR_WORKSPACE,
             "allreduce_init: the symmetric buffer must hold signals + staging + scratch (%zu bytes)",
             SIGNAL_BYTES + 2 * max_message_bytes);
  Context* c = new Context();
  c->rank = rank;
  c->world = world;
  cudaGetDevice(&c->device);
  for (int r = 0; r < world; r++) {
    B200_CHECK(buffer_ptrs[r] && ((uintptr_t)buffer_ptrs[r] & 1023) == 0, B200_ERR_INVALID,
               "allreduce_init: buffers must be non-null and 1024-byte aligned");
    c->regions[0].base[r] = reinterpret_cast<uint8_t*>(buffer_ptrs[r]);
  }
  c->regions[0].bytes = buffer_bytes;
  c->num_regions = 1;
  c->staging_off = SIGNAL_BYTES;
  c->staging_bytes = max_message_bytes;
  c->heap_bytes = buffer_bytes - SIGNAL_BYTES - 2 * max_message_bytes;
  c->mc_base = nullptr;
  *ctx_out = c;
  return B200_OK;
}

// NVSwitch multicast (NVLS) address of the symmetric buffer given to b200_allreduce_init: a virtual address bound
// (cuMulticastBindMem) to the SAME allocation on every rank, e.g. torch symmetric memory's multicast_ptr. With it
// b200_allreduce_run algo 3 and the fused kernel reduce inside the switch (multimem.ld_reduce / multimem.st).
// NULL switches back to peer loads / stores.
extern "C" int b200_allreduce_set_multicast(void* ctx, void* multicast_ptr) {
  B200_CHECK(ctx, B200_ERR_INVALID, "allreduce_set_multicast: null context");
  B200_CHECK(((uintptr_t)multicast_ptr & 1023) == 0, B200_ERR_INVALID,
             "allreduce_set_multicast: the multicast address must be 1024-byte aligned");
  reinterpret_cast<Context*>(ctx)->mc_base = reinterpret_cast<uint8_t*>(multicast_ptr);
  return B200_OK;
}
extern "C" int b200_allreduce_has_multicast(void* ctx) {
  return ctx && reinterpret_cast<Context*>(ctx)->mc_base ? 1 : 0;
}

// byte offset of the user heap inside the symmetric buffer, and its size
extern "C" int b200_allreduce_heap(void* ctx, size_t* offset, size_t* bytes) {
  B200_CHECK(ctx && offset && bytes, B200_ERR_INVALID, "allreduce_heap: null pointer");
  Context* c = reinterpret_cast<Context*>(ctx);
  *offset = c->staging_off + 2 * c->staging_bytes;
  *bytes = c->heap_bytes;
  return B200_OK;
}

// register one more symmetric buffer (peer pointers already opened with b200_ipc_open)
extern "C" int b200_allreduce_register(void* ctx, void* const* buffer_ptrs, size_t bytes) {
  B200_CHECK(ctx && buffer_ptrs && bytes > 0, B200_ERR_INVALID, "allreduce_register: bad arguments");
  Context* c = reinterpret_cast<Context*>(ctx);
  B200_CHECK(c->num_regions < MAX_REGIONS, B200_ERR_UNSUPPORTED, "allreduce_register: too many buffers");
  Region& g = c->regions[c->num_regions];
  for (int r = 0; r < c->world; r++) {
    B200_CHECK(buffer_ptrs[r] && ((uintptr_t)buffer_ptrs[r] & 15) == 0, B200_ERR_INVALID,
               "allreduce_register: buffers must be non-null and 16-byte aligned");
    g.base[r] = reinterpret_cast<uint8_t*>(buffer_ptrs[r]);
  }
  g.bytes = bytes;
  c->num_regions++;
  return B200_OK;
}

extern "C" int b200_allreduce_destroy(void* ctx) {
  delete reinterpret_cast<Context*>(ctx);
  return B200_OK;
}

// ---------------------------------------------------------------- collectives
// algo bit 2 (value 4) = skip the end barrier of the one-shot kernel (caller alternates input buffers).
// algo (low 2 bits): 0 = auto (thresholds of custom_all_reduce.cuh: one-shot for world 2, < 512 KiB at world 4,
// < 256 KiB at world 6/8; two-shot above), 1 = one-shot, 2 = two-shot, 3 = NVSwitch multimem (in-switch reduction).
extern "C" int b200_allreduce_run(void* ctx, const void* in, void* out, int64_t numel, int dtype, int algo,
                                  b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(ctx, B200_ERR_INVALID, "all_reduce: null context");
  Context* c = reinterpret_cast<Context*>(ctx);
  B200_CHECK(dtype == B200_F32 || dtype == B200_F16 || dtype == B200_BF16, B200_ERR_INVALID,
             "all_reduce: dtype must be float32 / float16 / bfloat16");
  B200_CHECK(numel >= 0, B200_ERR_INVALID, "all_reduce: bad size");
  if (numel == 0) return B200_OK;
  B200_CHECK(in && out, B200_ERR_INVALID, "all_reduce: null pointer");
  const size_t bytes = (size_t)numel * dtype_size(dtype);
  B200_CHECK(bytes % 16 == 0, B200_ERR_INVALID,
             "custom allreduce currently requires input length to be multiple of %d", 16 / dtype_size(dtype));
  B200_CHECK(((uintptr_t)in & 15) == 0 && ((uintptr_t)out & 15) == 0, B200_ERR_INVALID,
             "all_reduce: input / output must be 16-byte aligned");
  B200_CHECK(bytes <= c->staging_bytes, B200_ERR_WORKSPACE,
             "all_reduce: message of %zu bytes exceeds max size %zu", bytes, c->staging_bytes);
  Peers pe;
  const uint8_t* stage_src;
  int rc = resolve_input(c, in, bytes, &pe, &stage_src, "all_reduce");
  if (rc) return rc;
  const int skip_end = (algo & 4) ? 1 : 0;
  algo &= 3;
  const int64_t packs_mc = (int64_t)(bytes / 16);
  if (algo == 3) {
    // in-switch reduction; the output is broadcast with multimem.st when it lives in the symmetric buffer too
    B200_CHECK(c->mc_base && pe.mc_in, B200_ERR_UNSUPPORTED,
               "all_reduce: algo 3 (NVSwitch multimem) needs a multicast mapping (b200_allreduce_set_multicast) and an "
               "input inside the first symmetric buffer");
    size_t off_o;
    uint8_t* mc_out = nullptr;
    if (find_region(c, out, bytes, &off_o) == 0 && off_o >= c->staging_off + 2 * c->staging_bytes)
      mc_out = c->mc_base + off_o;
    if (dtype == B200_F32) return dispatch_mc<float>(c->world, pe, out, mc_out, stage_src, c->rank, packs_mc, stream);
    if (dtype == B200_F16) return dispatch_mc<__half>(c->world, pe, out, mc_out, stage_src, c->rank, packs_mc, stream);
    return dispatch_mc<__nv_bfloat16>(c->world, pe, out, mc_out, stage_src, c->rank, packs_mc, stream);
  }
  bool two_shot;
  if (algo == 1) two_shot = false;
  else if (algo == 2) two_shot = true;
  else two_shot = !(c->world == 2 || (c->world <= 4 && bytes < 512 * 1024) || (c->world <= 8 && bytes < 256 * 1024));
  const int64_t packs = (int64_t)(bytes / 16);
  if (dtype == B200_F32) return dispatch_ar<float>(c->world, pe, out, stage_src, c->rank, packs, two_shot, skip_end, stream);
  if (dtype == B200_F16) return dispatch_ar<__half>(c->world, pe, out, stage_src, c->rank, packs, two_shot, skip_end, stream);
  return dispatch_ar<__nv_bfloat16>(c->world, pe, out, stage_src, c->rank, packs, two_shot, skip_end, stream);
}

// out (all ranks, [M, hidden], must live in the symmetric heap at the same offset on every rank):
//   e4m3(static quant) if scale != NULL else dtype. residual: this rank's [rows_per_rank, hidden] rows
//   (rows_per_rank = ceil(M / world)), updated in place; NULL = no residual add.
extern "C" int b200_allreduce_fused_add_rms_norm(void* ctx, void* out, const void* in, void* residual,
                                                 const void* weight, const float* scale, float eps,
                                                 int64_t num_tokens, int hidden, int dtype,
                                                 b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(ctx, B200_ERR_INVALID, "allreduce_fused_add_rms_norm: null context");
  Context* c = reinterpret_cast<Context*>(ctx);
  B200_CHECK(dtype == B200_F16 || dtype == B200_BF16, B200_ERR_INVALID,
             "allreduce_fused_add_rms_norm: dtype must be float16 / bfloat16");
  B200_CHECK(num_tokens >= 0 && num_tokens < (1 << 30) && hidden > 0 && hidden % 8 == 0 && hidden <= 16384,
             B200_ERR_INVALID, "allreduce_fused_add_rms_norm: hidden must be a multiple of 8 and <= 16384");
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(in && out && weight, B200_ERR_INVALID, "allreduce_fused_add_rms_norm: null pointer");
  B200_CHECK(((uintptr_t)in & 15) == 0 && ((uintptr_t)out & 15) == 0 && ((uintptr_t)weight & 15) == 0 &&
                 ((uintptr_t)residual & 15) == 0,
             B200_ERR_INVALID, "allreduce_fused_add_rms_norm: pointers must be 16-byte aligned");
  const size_t bytes = (size_t)num_tokens * hidden * 2;
  Peers pe;
  const uint8_t* stage_src;
  int rc = resolve_input(c, in, bytes, &pe, &stage_src, "allreduce_fused_add_rms_norm");
  if (rc) return rc;
  const size_t out_bytes = (size_t)num_tokens * hidden * (scale ? 1 : 2);
  size_t off;
  const int reg = find_region(c, out, out_bytes, &off);
  B200_CHECK(reg >= 0, B200_ERR_INVALID,
             "allreduce_fused_add_rms_norm: out must live in the symmetric heap (it is written by the peers)");
  FusedParams fp;
  for (int r = 0; r < c->world; r++) fp.out[r] = c->regions[reg].base[r] + off;
  static const int mc_off = [] { const char* e = getenv("B200_AR_MULTIMEM"); return e && atoi(e) == 0; }();
  fp.mc_out = (reg == 0 && c->mc_base && !mc_off && hidden % (16 / 2) == 0) ? c->mc_base + off : nullptr;
  fp.residual = residual;
  fp.weight = weight;
  fp.scale = scale;
  fp.eps = eps;
  fp.M = (int)num_tokens;
  fp.hidden = hidden;
  fp.rows_per_rank = (int)((num_tokens + c->world - 1) / c->world);
  if (dtype == B200_BF16)
    return scale ? dispatch_fused<__nv_bfloat16, true>(c->world, pe, fp, stage_src, c->rank, stream)
                 : dispatch_fused<__nv_bfloat16, false>(c->world, pe, fp, stage_src, c->rank, stream);
  return scale ? dispatch_fused<__half, true>(c->world, pe, fp, stage_src, c->rank, stream)
               : dispatch_fused<__half, false>(c->world, pe, fp, stage_src, c->rank, stream);
}

// ---------------------------------------------------------------- DCP attention merge (csrc/dcp_merge.cuh)
// cp_lse_ag_out_rs / cp_lse_ag_out_ar (vllm/v1/attention/ops/common.py:185-258) in one NVLink kernel.
// attn_out [B, H, D] (dtype) and attn_lse [B, H] (f32): this rank's partial attention over its shard of the
// context, BOTH inside a registered symmetric region at the same offset on every rank (peers read them in place).
// mode 0 (reduce-scatter over heads): out [B, H / world, D], lse_out [B, H / world] = heads [rank*H/world, ...);
// mode 1 (all-reduce): out [B, H, D], lse_out [B, H]. out / lse_out are local (any device memory), lse_out nullable.
extern "C" int b200_dcp_lse_merge(void* ctx, void* out, float* lse_out, const void* attn_out, const float* attn_lse,
                                  int64_t num_tokens, int num_heads, int head_size, int dtype, int mode,
                                  int is_lse_base_on_e, b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(ctx, B200_ERR_INVALID, "dcp_lse_merge: null context");
  Context* c = reinterpret_cast<Context*>(ctx);
  B200_CHECK(dtype == B200_F32 || dtype == B200_F16 || dtype == B200_BF16, B200_ERR_INVALID,
             "dcp_lse_merge: dtype must be float32 / float16 / bfloat16");
  B200_CHECK(mode == 0 || mode == 1, B200_ERR_INVALID, "dcp_lse_merge: mode must be 0 (reduce-scatter) or 1 (all-reduce)");
  B200_CHECK(num_tokens >= 0 && num_tokens < (1 << 24) && num_heads > 0 && head_size > 0, B200_ERR_INVALID,
             "dcp_lse_merge: bad sizes");
  B200_CHECK((head_size * dtype_size(dtype)) % 16 == 0, B200_ERR_INVALID,
             "dcp_lse_merge: head_size rows must be multiples of 16 bytes");
  B200_CHECK(mode == 1 || num_heads % c->world == 0, B200_ERR_INVALID,
             "dcp_lse_merge: num_heads (%d) must divide by the group size (%d) for reduce-scatter", num_heads, c->world);
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(out && attn_out && attn_lse, B200_ERR_INVALID, "dcp_lse_merge: null pointer");
  B200_CHECK(((uintptr_t)out & 15) == 0 && ((uintptr_t)attn_out & 15) == 0 && ((uintptr_t)attn_lse & 3) == 0,
             B200_ERR_INVALID, "dcp_lse_merge: out / attn_out must be 16-byte aligned");
  const size_t out_bytes = (size_t)num_tokens * num_heads * head_size * dtype_size(dtype);
  const size_t lse_bytes = (size_t)num_tokens * num_heads * 4;
  Peers pe;
  DcpParams dp;
  size_t off_o, off_l;
  const int reg_o = find_region(c, attn_out, out_bytes, &off_o);
  const int reg_l = find_region(c, attn_lse, lse_bytes, &off_l);
  B200_CHECK(reg_o >= 0 && reg_l >= 0, B200_ERR_INVALID,
             "dcp_lse_merge: attn_out and attn_lse must live in the symmetric heap (peers read them in place)");
  pe.pdl_mode = 2;   // the DCP kernel is launched plainly
  pe.spin_ns = spin_timeout_ns();
  for (int r = 0; r < c->world; r++) {
    pe.sig[r] = reinterpret_cast<Signal*>(c->regions[0].base[r]);
    pe.tmp[r] = nullptr;
    pe.in[r] = c->regions[reg_o].base[r] + off_o;
    dp.lse[r] = reinterpret_cast<const float*>(c->regions[reg_l].base[r] + off_l);
  }
  dp.out = out;
  dp.lse_out = lse_out;
  dp.B = (int)num_tokens;
  dp.H = num_heads;
  dp.D = head_size;
  dp.Hs = mode == 0 ? num_heads / c->world : num_heads;
  dp.head_lo = mode == 0 ? c->rank * dp.Hs : 0;
  dp.base_e = is_lse_base_on_e ? 1 : 0;
  const int64_t packs = (int64_t)dp.B * dp.Hs * (head_size * dtype_size(dtype) / 16);
  int64_t blocks = (packs + 255) / 256;
  const int cap = sm_count() < MAX_BLOCKS ? sm_count() : MAX_BLOCKS;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
#define B200_DCP(T, N) dcp_lse_merge_kernel<T, N><<<(int)blocks, 256, 0, stream>>>(pe, dp, c->rank)
#define B200_DCP_T(T)                                                                      \
  switch (c->world) {                                                                      \
    case 2: B200_DCP(T, 2); break;                                                         \
    case 4: B200_DCP(T, 4); break;                                                         \
    case 6: B200_DCP(T, 6); break;                                                         \
    case 8: B200_DCP(T, 8); break;                                                         \
    default:                                                                               \
      set_last_error("dcp_lse_merge supports 2, 4, 6 or 8 ranks, got %d", c->world);       \
      return B200_ERR_UNSUPPORTED;                                                         \
  }
  if (dtype == B200_F32) { B200_DCP_T(float) }
  else if (dtype == B200_F16) { B200_DCP_T(__half) }
  else { B200_DCP_T(__nv_bfloat16) }
#undef B200_DCP_T
#undef B200_DCP
  B200_LAUNCH_CHECK();
  return B200_OK;
}
