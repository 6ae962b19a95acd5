// Internal declarations shared by the host runtime and the CUDA translation units.
#ifndef B200_INTERNAL_H
#define B200_INTERNAL_H

#include <cstddef>
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace b200 {

// records the message and returns code (thread-local last-error string)
int fail(int code, const char* msg);
int failf(int code, const char* fmt, ...);
int owner_of(int64_t gid, int64_t n_global, int nranks);

}  // namespace b200

#ifdef __CUDACC__
#define B200_HD __host__ __device__
#else
#define B200_HD
#endif

#endif
