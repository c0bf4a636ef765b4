// oracle/dxm_shim/DirectXPackedVector.h -- TEST INFRASTRUCTURE.  XMCOLOR and its load/store live in DirectXMath.h of this shim.
#pragma once
#include "DirectXMath.h"
