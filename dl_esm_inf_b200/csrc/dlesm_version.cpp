// dlesm_version.cpp -- the library's identity.  dl_esm_inf_b200/build.py compiles this unit on every
// build with -DDLESM_SOURCE_HASH="<sha256 of every source, header and flag>", so that
// __graft_entry__.build() can tell a library built from the sources on disk from a stale or
// foreign one (it scans the .so for the marker, and dlesm_source_hash() returns the same string).
#include "../../include/dlesm_b200.h"

#ifndef DLESM_SOURCE_HASH
#define DLESM_SOURCE_HASH "unhashed"
#endif

namespace {
// referenced by dlesm_source_hash(), so the marker string stays in .rodata
const char kMarker[] = "DLESM_SRC_HASH=" DLESM_SOURCE_HASH;
}  // namespace

extern "C" const char *dlesm_source_hash(void) { return kMarker + 15; }
extern "C" const char *dlesm_version(void)
{
    return "dlesm_b200 0.2 (sm_100a, FP64; default arithmetic fmad=false, contracted twin of the fused sweeps "
           "selectable) src " DLESM_SOURCE_HASH;
}
