/* oracle shim -- TEST INFRASTRUCTURE ONLY.
 * The reference (src/c++/pipes/impl/HadoopPipes.cc:421-429) uses the
 * OpenSSL 0.9/1.0 stack-allocated HMAC_CTX API, which OpenSSL 3 removed
 * (HMAC_CTX is opaque now).  This wrapper keeps that call shape compiling on
 * top of the heap-allocated API.  Only reached when a job-token password is
 * set, i.e. never in command-file replay mode. */
#ifndef HGM_ORACLE_SHIM_HMAC_H
#define HGM_ORACLE_SHIM_HMAC_H
#define OPENSSL_SUPPRESS_DEPRECATED 1
#include_next <openssl/hmac.h>
#include <openssl/evp.h>
typedef struct { HMAC_CTX *p; } hgm_shim_hmac_ctx;
static inline int hgm_shim_hmac_init(hgm_shim_hmac_ctx *c, const void *key, int len, const EVP_MD *md) {
  c->p = HMAC_CTX_new();
  return HMAC_Init_ex(c->p, key, len, md, NULL);
}
static inline int hgm_shim_hmac_update(hgm_shim_hmac_ctx *c, const unsigned char *d, size_t n) {
  return HMAC_Update(c->p, d, n);
}
static inline int hgm_shim_hmac_final(hgm_shim_hmac_ctx *c, unsigned char *md, unsigned int *len) {
  return HMAC_Final(c->p, md, len);
}
static inline void hgm_shim_hmac_cleanup(hgm_shim_hmac_ctx *c) { HMAC_CTX_free(c->p); c->p = NULL; }
#define HMAC_CTX     hgm_shim_hmac_ctx
#define HMAC_Init    hgm_shim_hmac_init
#define HMAC_Update  hgm_shim_hmac_update
#define HMAC_Final   hgm_shim_hmac_final
#define HMAC_cleanup hgm_shim_hmac_cleanup
#endif
