// C-ABI: library-level entry points (version, error string, launch counter).
#include "common.cuh"

extern "C" int peft_version(void) { return 100; }

extern "C" void peft_last_error(char* buf, size_t n) {
  if (buf == nullptr || n == 0) return;
  strncpy(buf, peft::tls_error(), n - 1);
  buf[n - 1] = '\0';
}

extern "C" unsigned long long peft_launch_count(void) { return peft::launch_counter().load(); }
extern "C" unsigned long long peft_kernel_count(int kind) {
  return (kind >= 0 && kind < peft::KIND_COUNT) ? peft::kind_counter(kind).load() : 0ull;
}
