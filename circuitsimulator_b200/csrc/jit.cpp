// jit.cpp — see jit.hpp.
#include "jit.hpp"

#include <dlfcn.h>

#include <cstring>
#include <mutex>

namespace csim {
namespace {

// the handful of NVRTC entry points we need, resolved at run time
struct Nvrtc {
    void* so = nullptr;
    int (*CreateProgram)(void**, const char*, const char*, int, const char* const*, const char* const*) = nullptr;
    int (*CompileProgram)(void*, int, const char* const*) = nullptr;
    int (*GetProgramLogSize)(void*, size_t*) = nullptr;
    int (*GetProgramLog)(void*, char*) = nullptr;
    int (*GetCUBINSize)(void*, size_t*) = nullptr;
    int (*GetCUBIN)(void*, char*) = nullptr;
    int (*DestroyProgram)(void**) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool ok = false;
};

Nvrtc& nvrtc() {
    static Nvrtc n;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = {"libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so.12",
                               "/usr/local/cuda/lib64/libnvrtc.so"};
        for (const char* nm : names) {
            n.so = dlopen(nm, RTLD_NOW | RTLD_LOCAL);
            if (n.so) break;
        }
        if (!n.so) return;
        auto sym = [&](const char* s) { return dlsym(n.so, s); };
        n.CreateProgram = reinterpret_cast<decltype(n.CreateProgram)>(sym("nvrtcCreateProgram"));
        n.CompileProgram = reinterpret_cast<decltype(n.CompileProgram)>(sym("nvrtcCompileProgram"));
        n.GetProgramLogSize = reinterpret_cast<decltype(n.GetProgramLogSize)>(sym("nvrtcGetProgramLogSize"));
        n.GetProgramLog = reinterpret_cast<decltype(n.GetProgramLog)>(sym("nvrtcGetProgramLog"));
        n.GetCUBINSize = reinterpret_cast<decltype(n.GetCUBINSize)>(sym("nvrtcGetCUBINSize"));
        n.GetCUBIN = reinterpret_cast<decltype(n.GetCUBIN)>(sym("nvrtcGetCUBIN"));
        n.DestroyProgram = reinterpret_cast<decltype(n.DestroyProgram)>(sym("nvrtcDestroyProgram"));
        n.GetErrorString = reinterpret_cast<decltype(n.GetErrorString)>(sym("nvrtcGetErrorString"));
        n.ok = n.CreateProgram && n.CompileProgram && n.GetProgramLogSize && n.GetProgramLog && n.GetCUBINSize &&
               n.GetCUBIN && n.DestroyProgram;
    });
    return n;
}

// NVRTC has no host standard library: the device headers only need the fixed-width typedefs.
const char kStdintShim[] =
    "#pragma once\n"
    "typedef signed char int8_t; typedef unsigned char uint8_t; typedef short int16_t; typedef unsigned short uint16_t;\n"
    "typedef int int32_t; typedef unsigned int uint32_t; typedef long long int64_t; typedef unsigned long long uint64_t;\n";

}  // namespace

std::string jit_library_dir() {
    Dl_info info;
    if (dladdr(reinterpret_cast<void*>(&jit_library_dir), &info) && info.dli_fname) {
        std::string p(info.dli_fname);
        auto slash = p.find_last_of('/');
        return slash == std::string::npos ? std::string(".") : p.substr(0, slash);
    }
    return ".";
}

bool jit_compile(const std::string& source, const std::string& name, const std::string& include_dir,
                 std::vector<char>& cubin, std::string& log) {
    Nvrtc& n = nvrtc();
    if (!n.ok) {
        log = "libnvrtc not available";
        return false;
    }
    void* prog = nullptr;
    const char* hdr_src[] = {kStdintShim};
    const char* hdr_name[] = {"stdint.h"};
    int rc = n.CreateProgram(&prog, source.c_str(), name.c_str(), 1, hdr_src, hdr_name);
    if (rc != 0) {
        log = "nvrtcCreateProgram failed";
        return false;
    }
    const std::string inc = "-I" + include_dir;
    // (177: generated code declares every static value, used or not)
    const char* opts[] = {"--gpu-architecture=sm_100a", "--std=c++17", "--fmad=false", "-lineinfo", "--diag-suppress=177",
                          inc.c_str()};
    rc = n.CompileProgram(prog, 6, opts);
    size_t ls = 0;
    n.GetProgramLogSize(prog, &ls);
    if (ls > 1) {
        log.resize(ls);
        n.GetProgramLog(prog, &log[0]);
    }
    bool ok = rc == 0;
    if (ok) {
        size_t cs = 0;
        ok = n.GetCUBINSize(prog, &cs) == 0 && cs > 0;
        if (ok) {
            cubin.resize(cs);
            ok = n.GetCUBIN(prog, cubin.data()) == 0;
        }
    } else if (n.GetErrorString) {
        log = std::string(n.GetErrorString(rc)) + "\n" + log;
    }
    n.DestroyProgram(&prog);
    return ok;
}

cudaError_t jit_load(const std::vector<char>& cubin, const char* entry, JitKernel& out) {
    cudaError_t e = cudaLibraryLoadData(&out.lib, cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0);
    if (e != cudaSuccess) return e;
    e = cudaLibraryGetKernel(&out.kernel, out.lib, entry);
    if (e != cudaSuccess) {
        cudaLibraryUnload(out.lib);
        out.lib = nullptr;
    }
    return e;
}

void jit_unload(JitKernel& k) {
    if (k.lib) cudaLibraryUnload(k.lib);
    k.lib = nullptr;
    k.kernel = nullptr;
}

}  // namespace csim
