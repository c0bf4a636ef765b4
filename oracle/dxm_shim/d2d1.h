// oracle/dxm_shim/d2d1.h -- TEST INFRASTRUCTURE.  Empty: Direct2D is only used by app.cpp / spheres-app.cpp, which are not compiled.
#pragma once
