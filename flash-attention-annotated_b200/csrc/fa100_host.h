// fa100_host.h -- host-side helpers shared by the translation units of libfa100.so.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include "../../include/fa100.h"

namespace fa100 {

int set_error(int code, const char* fmt, ...);
const char* last_error();

#define FA_CHECK_ARG(cond, ...)                                  \
  do {                                                           \
    if (!(cond)) return fa100::set_error(FA100_ERR_INVALID, __VA_ARGS__); \
  } while (0)
#define FA_UNSUPPORTED(cond, ...)                                \
  do {                                                           \
    if (cond) return fa100::set_error(FA100_ERR_UNSUPPORTED, __VA_ARGS__); \
  } while (0)
#define FA_CUDA(expr)                                                                                   \
  do {                                                                                                  \
    cudaError_t e_ = (expr);                                                                            \
    if (e_ != cudaSuccess)                                                                              \
      return fa100::set_error(FA100_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

// cuTensorMapEncodeTiled through the runtime's driver entry point (no -lcuda link dependency).
// dims/strides innermost first; strides in bytes for dims 1..rank-1; 128-byte swizzle, zero OOB fill.
int encode_tensor_map(CUtensorMap* out, int dtype, int rank, const void* base, const uint64_t* dims,
                      const uint64_t* strides_bytes, const uint32_t* box);

int device_sm_count(int* out);

// Launch with programmatic stream serialization (PDL): the kernel may start while the previous kernel of the stream is
// still draining; every fa100 kernel executes griddepcontrol.wait before its first global-memory access (fa100_ptx.cuh).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, int cluster_x, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int n = 0;
#ifndef FA100_NO_PDL  // A/B switch of the build (tools/build_variant.sh): plain stream-ordered launches
  attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[n].val.programmaticStreamSerializationAllowed = 1;
  ++n;
#endif
  if (cluster_x > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = cluster_x; attr[n].val.clusterDim.y = 1; attr[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg.attrs = attr; cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

int launch_fwd(const Fa100FwdParams* p, const Fa100Plan* plan, cudaStream_t stream);
int launch_combine(const Fa100CombineParams* p, cudaStream_t stream);
int launch_prepare_scheduler(const Fa100FwdParams* p, const Fa100Plan* plan, int32_t* metadata, int64_t words, cudaStream_t stream);
int64_t scheduler_metadata_words(int b);
int make_plan(const Fa100Shape* s, Fa100Plan* out);

}  // namespace fa100
