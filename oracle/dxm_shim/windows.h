// oracle/dxm_shim/windows.h -- TEST INFRASTRUCTURE.  Stand-in so the reference's stdafx.h (src/common-lib/stdafx.h:24)
// preprocesses on Linux.  The reference files compiled by oracle/Makefile use nothing from the Win32 API; the only
// MSVC-ism they contain is __declspec(align(16)) (ray-tracing.h:5,15,55), which XMVECTOR's own alignment makes redundant.
#pragma once
#include <functional>   // light.h uses std::function without including <functional> (MSVC pulls it in transitively)
#define __declspec(x)
