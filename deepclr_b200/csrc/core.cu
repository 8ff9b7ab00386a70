// Error plumbing and identification for libdeepclr_b200.
#include "common.cuh"
#include <string.h>

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

}  // namespace dclr

extern "C" const char *dclr_version(void) { return "deepclr_b200 0.1 (sm_100a)"; }
extern "C" const char *dclr_last_error(void) { return dclr::err_buf(); }
