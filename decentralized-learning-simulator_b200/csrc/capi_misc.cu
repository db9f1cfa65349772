#include "common.cuh"
namespace dls { int64_t g_launch_count = 0; }
extern "C" int dls_version(void) { return 100; }
extern "C" const char *dls_build_info(void) {
    return "dls_b200 sm_100a nvcc " __DATE__ " " __TIME__;
}
extern "C" int64_t dls_launch_count(void) { return dls::g_launch_count; }
