// ABI bookkeeping shared by every translation unit of libllmtrain_b200.so.
#include "host.cuh"
#include "../../include/llmtrain_b200.h"

extern "C" const char* llmtrain_b200_last_error(void) { return b200::last_error_slot().c_str(); }
extern "C" int llmtrain_b200_abi_version(void) { return 1000; }
extern "C" long long llmtrain_b200_launch_count(void) {
  return b200::launch_counter().load(std::memory_order_relaxed);
}
extern "C" void llmtrain_b200_set_deterministic(int on) {
  b200::deterministic_flag().store(on != 0 ? 1 : 0, std::memory_order_relaxed);
}
extern "C" int llmtrain_b200_get_deterministic(void) { return b200::deterministic() ? 1 : 0; }
