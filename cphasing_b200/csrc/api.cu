// Error reporting, version and launch accounting of libcph.
#include "common.cuh"
#include <string.h>

unsigned long long g_cph_launches = 0;

namespace {
thread_local char g_err[512] = "";
}

void cph_set_error(const char* file, int line, const char* what) {
    const char* base = strrchr(file, '/');
    snprintf(g_err, sizeof(g_err), "%s:%d: %s", base ? base + 1 : file, line, what);
}

extern "C" {
int cph_version(void) { return 100; }
const char* cph_last_error(void) { return g_err; }
uint64_t cph_launch_count(void) { return g_cph_launches; }
}
