// TEST INFRASTRUCTURE ONLY (oracle). Shadowing shim for Common/odprintf.h:
// the release build of the reference compiles odprintf to a no-op (odprintf.h:41-45).
#ifndef FS3_ORACLE_SHIM_ODPRINTF_H
#define FS3_ORACLE_SHIM_ODPRINTF_H
static inline void odprintf(const char*, ...) {}
#endif
