// TEST INFRASTRUCTURE ONLY (oracle). Shadowing shim for the reference's
// Common/MathHelper.h so that /root/reference/FluidSimulationDemo/Simulation3D.cpp
// compiles headless on Linux without <Windows.h>/<DirectXMath.h>.
// Only the two templates FluidSim3 uses are provided, with the semantics of
// Common/MathHelper.h:43-53 (Lerp = a + (b-a)*t, Clamp = x<lo?lo:(x>hi?hi:x)),
// plus PODs standing in for DirectX::XMFLOAT2/XMFLOAT3.
#ifndef FS3_ORACLE_SHIM_MATHHELPER_H
#define FS3_ORACLE_SHIM_MATHHELPER_H
#include <cmath>
#include <cstring>
#include <cassert>
#include <cstdlib>
#include <cfloat>

namespace DirectX {
struct XMFLOAT2 { float x, y; XMFLOAT2() : x(0), y(0) {} XMFLOAT2(float a, float b) : x(a), y(b) {} };
struct XMFLOAT3 { float x, y, z; XMFLOAT3() : x(0), y(0), z(0) {} XMFLOAT3(float a, float b, float c) : x(a), y(b), z(c) {} };
}

class MathHelper {
public:
	template <typename T> static T Min(const T& a, const T& b) { return a < b ? a : b; }
	template <typename T> static T Max(const T& a, const T& b) { return a > b ? a : b; }
	template <typename T> static T Lerp(const T& a, const T& b, float t) { return a + (b - a) * t; }
	template <typename T> static T Clamp(const T& x, const T& lo, const T& hi) { return x < lo ? lo : (x > hi ? hi : x); }
};
#endif
