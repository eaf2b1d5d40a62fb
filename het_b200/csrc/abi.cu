// ABI bookkeeping for libhetb200: version, build string, launch counter.
#include "common.cuh"

namespace hetb {
int64_t g_launch_count = 0;
}

extern "C" int hetb_abi_version(void) { return 1; }

extern "C" const char *hetb_build_info(void) {
  return "hetb200 (B200-native HET/Hector hot path) | arch sm_100a | nvcc "
#define HETB_STR2(x) #x
#define HETB_STR(x) HETB_STR2(x)
      HETB_STR(__CUDACC_VER_MAJOR__) "." HETB_STR(__CUDACC_VER_MINOR__) "." HETB_STR(
          __CUDACC_VER_BUILD__) " | built " __DATE__ " " __TIME__;
}

extern "C" int64_t hetb_launch_count(void) { return hetb::g_launch_count; }
