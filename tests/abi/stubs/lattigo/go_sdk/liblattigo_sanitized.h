/* TEST STUB for the cgo-generated header the reference's fhe_lib_v2.h includes (fhe_ops_lib/lattigo is an absent
 * submodule).  Only what the declarations in fhe_lib_v2.h need to PARSE; nothing here is ever called. */
#pragma once
#include <stdint.h>
#include <stddef.h>
typedef long long GoInt64;
typedef unsigned long long GoUint64;
typedef int GoInt32;
typedef unsigned int GoUint32;
typedef GoInt64 GoInt;
typedef GoUint64 GoUint;
typedef double GoFloat64;
typedef unsigned char GoUint8;
typedef size_t GoUintptr;
#ifdef __cplusplus
extern "C" {
#endif
void ReleaseHandle(uint64_t handle);
#ifdef __cplusplus
}
#endif
