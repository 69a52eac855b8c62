// TEST INFRASTRUCTURE ONLY: force-included (with -DPARS_DEBUG_H) in place of the reference's PS_FrameWork/include/_parsDebug.h,
// whose second half includes GL/gl.h for glGetError helpers (no graphics stack on this box). Only the first half is used by the
// scene API (PS_BlobTree/include/CAdaptiveUniformGrid3d.h:10): ASSERT and parsDebugInfo.
#ifndef PSGPU_ORACLE_PARSDEBUG_MIN_H
#define PSGPU_ORACLE_PARSDEBUG_MIN_H
#include <string>
#include <cstdio>
#include <cstdarg>
using namespace std;
#ifndef ASSERT
#define ASSERT(x)
#endif
static inline void parsDebugInfo(char* str, ...) { va_list a; va_start(a, str); vfprintf(stderr, str, a); va_end(a); }
#endif
