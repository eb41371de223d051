// Error-string plumbing shared by all launchers.
#include "common.cuh"
#include <string.h>

static thread_local char g_err[512] = "ok";

void dsb200_set_error(const char* msg)
{
    strncpy(g_err, msg, sizeof(g_err) - 1);
    g_err[sizeof(g_err) - 1] = 0;
}

extern "C" const char* dsb200_last_error() { return g_err; }

extern "C" int dsb200_version() { return 1; }
