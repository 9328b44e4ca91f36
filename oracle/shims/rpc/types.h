/* oracle shim -- TEST INFRASTRUCTURE ONLY.
 * glibc in this image ships no Sun RPC headers; the reference's
 * src/c++/utils/impl/SerialUtils.cc:22-23 includes <rpc/types.h> and
 * <rpc/xdr.h> only to (de)serialise one float (SerialUtils.cc:246-262).
 * This header plus rpc/xdr.h give it a minimal big-endian IEEE-754 stand-in. */
#ifndef HGM_ORACLE_SHIM_RPC_TYPES_H
#define HGM_ORACLE_SHIM_RPC_TYPES_H
#include <stdint.h>
typedef int bool_t;
#endif
