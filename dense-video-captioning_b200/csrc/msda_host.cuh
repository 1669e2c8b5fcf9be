// msda_host.cuh -- host-side helpers shared by the translation units of libmsda_sm100.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/msda.h"

namespace msda {
namespace host {

// Diagnostics only (msda_last_launch_count / msda_last_launch_path): per-thread, reset by every entry point.
extern thread_local int g_launches;
extern thread_local int g_path;

// Tuning knobs are compiled in only with -DMSDA_TUNING (development builds, tools/); the shipped library reads no
// environment variable.
inline int env_int(const char *name, int dflt)
{
#ifdef MSDA_TUNING
    const char *e = getenv(name);
    return e ? atoi(e) : dflt;
#else
    (void)name;
    return dflt;
#endif
}

// SM count of the current device (grids are sized in multiples of it): queried once per device ordinal, an
// immutable device property.
inline int num_sms()
{
    static int cache[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int v = cache[dev];
    if (v == 0) {
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        cache[dev] = v;
    }
    return v;
}

inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace host
}  // namespace msda
