// jit.h — run-time compiled fused aggregate kernels (jit.cu)
#pragma once
#include <cuda_runtime.h>

#include <memory>
#include <string>

namespace b2 {

struct JitKernels {  // CUfunction handles of one compiled shape
  void* tma_narrow = nullptr;  // agg_tma_kernel<S, true>  (null when the key shape has no narrow form)
  void* tma_wide = nullptr;    // agg_tma_kernel<S, false>
  void* small = nullptr;       // agg_small_kernel<S>
  void* global = nullptr;      // agg_global_kernel<S>
  int static_smem_narrow = 0, static_smem_wide = 0;
  int per_sm_small = 1, per_sm_global = 1;
  bool narrow_ok = false;
  bool compiled_now = false;  // false: served from the disk cache
};

// signature text "p:...|k:...|a:..." -> template type texts; false when the shape is outside the fused vocabulary
bool jit_signature_types(const std::string& sig, std::string* pred_t, std::string* keys_t, std::string* args_t,
                         int* n_int_keys, int* n_str_keys);
// compiled kernels of a shape (compiles on first use; null + *why when it cannot)
std::shared_ptr<JitKernels> jit_fused_kernels(const std::string& sig, std::string* why);
// NVRTC compilation of a shape without loading it (needs no device): cubin bytes, or 0 and *why
size_t jit_compile_check(const std::string& sig, std::string* why);
void jit_launch(void* fn, int grid, int block, size_t dyn_smem, cudaStream_t stream, void** args);

}  // namespace b2
