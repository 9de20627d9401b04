// jit.hpp — NVRTC compilation (sm_100a) and loading of generated kernels.
// libnvrtc is dlopen()ed on first use so that the library itself loads on machines without
// the CUDA toolkit's run-time compiler; a missing NVRTC simply disables specialisation (the
// generic statically compiled kernels are used instead — same results, slower).
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

namespace csim {

// Compiles `source` to a cubin for sm_100a.  include_dir = directory holding the device headers
// (csrc/).  Returns false and fills `log` on failure.
bool jit_compile(const std::string& source, const std::string& name, const std::string& include_dir,
                 std::vector<char>& cubin, std::string& log);

// Directory of the shared object that contains this function (to locate csrc/ at run time).
std::string jit_library_dir();

struct JitKernel {
    cudaLibrary_t lib = nullptr;
    cudaKernel_t kernel = nullptr;
};
cudaError_t jit_load(const std::vector<char>& cubin, const char* entry, JitKernel& out);
void jit_unload(JitKernel& k);

}  // namespace csim
