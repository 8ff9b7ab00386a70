// Error plumbing and identification for libdeepclr_b200.
#include "common.cuh"
#include <string.h>
#include <atomic>

namespace dclr {

char *err_buf() {
    static thread_local char buf[512] = {0};
    return buf;
}

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err_buf(), 512, fmt, ap);
    va_end(ap);
    return code;
}

static std::atomic<int> g_fps_tie_rule{0};
int fps_tie_rule() { return g_fps_tie_rule.load(std::memory_order_relaxed); }

}  // namespace dclr

extern "C" int dclr_set_fps_tie_rule(int rule) {
    if (rule != 0 && rule != 1) return dclr::fail(DCLR_E_INVALID, "dclr_set_fps_tie_rule: rule must be 0 or 1 (got %d)", rule);
    dclr::g_fps_tie_rule.store(rule, std::memory_order_relaxed);
    return DCLR_OK;
}
extern "C" int dclr_get_fps_tie_rule(void) { return dclr::fps_tie_rule(); }

extern "C" const char *dclr_version(void) { return "deepclr_b200 0.1 (sm_100a)"; }
extern "C" const char *dclr_last_error(void) { return dclr::err_buf(); }
