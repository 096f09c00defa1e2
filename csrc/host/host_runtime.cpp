// Host-only runtime pieces of idc_models_b200 (no CUDA dependency): error string plumbing shared by
// the HDF5 subset, the checkpoint writer and the prefetch ring.
#include <cstdarg>
#include <cstdio>
#include <cstring>

#include "host_common.h"
#include "idcb200_host.h"

namespace idcbh {
static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace idcbh

extern "C" const char* idcbh_last_error(void) { return idcbh::g_err; }
extern "C" int idcbh_version(void) { return 100; }
