// Error plumbing, version and launch accounting shared by every entry point of libpdnn_b200.so.
#include "common.cuh"
#include "../include/pdnn_b200.h"

#include <atomic>
#include <stdarg.h>

static thread_local char g_err[1024] = "";
std::atomic<long long> g_pdnn_launches{0};

void pdnn_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* pdnn_last_error() { return g_err; }
extern "C" int pdnn_version() { return 100; }
extern "C" long long pdnn_launch_count() { return g_pdnn_launches.load(); }
extern "C" void pdnn_reset_launch_count() { g_pdnn_launches.store(0); }
// A CUDA-graph replay runs the library kernels that were captured into it without going through the host entry points:
// the owner of the graph adds the captured number per replay so that the counter keeps meaning "kernels run".
extern "C" void pdnn_add_launch_count(long long n) { g_pdnn_launches.fetch_add(n); }
