/*
 * rc4.h — drop-in for the reference's rc4.h (FreeBSD rc4, locally edited).
 * The vendored rc4.c is undefined behaviour as shipped (indices are never
 * reduced mod 256, rc4.c:58,68-71,88-93 — SURVEY finding 4); this
 * implementation provides its intended semantics (the upstream unsigned-byte
 * types named at rc4.c:36) = standard RC4 = arc4.c, with the state mapping
 * index1<->x, index2<->y, perm<->m.  Runs on the B200 engine.
 */
#ifndef BULKCIPHER_DROPIN_RC4_H
#define BULKCIPHER_DROPIN_RC4_H

/* ABI of rc4.h:43-47 (264 bytes) */
struct rc4_state {
    char perm[256];
    int index1;
    int index2;
};

#ifdef __cplusplus
extern "C" {
#endif

/* rc4.c:56-72 */
void rc4_init(struct rc4_state *const state, const char *key, int keylen);
/* rc4.c:80-99 — PRGA fused with the XOR; inbuf == outbuf allowed (:76) */
void rc4_crypt(struct rc4_state *const state, const char *inbuf, char *outbuf, int buflen);

#ifdef __cplusplus
}
#endif
#endif
