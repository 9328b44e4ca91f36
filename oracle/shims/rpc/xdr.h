/* oracle shim -- TEST INFRASTRUCTURE ONLY (see rpc/types.h).
 * XDR encodes a float as its 4 IEEE-754 bytes, most significant first. */
#ifndef HGM_ORACLE_SHIM_RPC_XDR_H
#define HGM_ORACLE_SHIM_RPC_XDR_H
#include <stdint.h>
#include <string.h>
#include "types.h"
enum xdr_op { XDR_ENCODE = 0, XDR_DECODE = 1 };
typedef struct { enum xdr_op op; char *mem; unsigned len; } XDR;
static inline void xdrmem_create(XDR *x, char *mem, unsigned len, enum xdr_op op) {
  x->op = op; x->mem = mem; x->len = len;
}
static inline bool_t xdr_float(XDR *x, float *f) {
  uint32_t bits;
  if (x->len < 4) return 0;
  if (x->op == XDR_ENCODE) {
    memcpy(&bits, f, 4);
    x->mem[0] = (char)(bits >> 24); x->mem[1] = (char)(bits >> 16);
    x->mem[2] = (char)(bits >> 8);  x->mem[3] = (char)bits;
  } else {
    bits = ((uint32_t)(unsigned char)x->mem[0] << 24) | ((uint32_t)(unsigned char)x->mem[1] << 16) |
           ((uint32_t)(unsigned char)x->mem[2] << 8)  |  (uint32_t)(unsigned char)x->mem[3];
    memcpy(f, &bits, 4);
  }
  return 1;
}
#endif
