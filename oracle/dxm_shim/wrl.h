// oracle/dxm_shim/wrl.h -- TEST INFRASTRUCTURE.  Empty: WRL ComPtr is only used by app.cpp / spheres-app.cpp, which are not compiled.
#pragma once
