/* Declarations-only stand-in for <gcrypt.h>, used ONLY to compile the unmodified
 * reference sources under /root/reference into oracle/_ref/ (test infrastructure).
 * The image ships libgcrypt.so.20 (runtime) but not its development header. The
 * prototypes and enum values below are libgcrypt's stable public ABI (1.8+);
 * nothing here is implementation. See SURVEY.md Appendix B. */
#ifndef ORACLE_GCRYPT_SHIM_H
#define ORACLE_GCRYPT_SHIM_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef unsigned int gpg_error_t;
typedef gpg_error_t gcry_error_t;
typedef struct gcry_md_handle *gcry_md_hd_t;
typedef struct gcry_cipher_handle *gcry_cipher_hd_t;

enum gcry_md_algos {
    GCRY_MD_MD5 = 1, GCRY_MD_RMD160 = 3, GCRY_MD_SHA256 = 8, GCRY_MD_SHA384 = 9,
    GCRY_MD_SHA512 = 10, GCRY_MD_CRC32 = 302, GCRY_MD_SHA3_256 = 313,
    GCRY_MD_SHA3_512 = 315, GCRY_MD_SHAKE128 = 316, GCRY_MD_SHAKE256 = 317
};
enum gcry_md_flags { GCRY_MD_FLAG_SECURE = 1 };
enum gcry_cipher_algos { GCRY_CIPHER_AES128 = 7, GCRY_CIPHER_AES256 = 9 };
enum gcry_cipher_modes { GCRY_CIPHER_MODE_CBC = 3 };
enum gcry_cipher_flags { GCRY_CIPHER_SECURE = 1, GCRY_CIPHER_CBC_CTS = 4 };
enum gcry_kdf_algos { GCRY_KDF_SCRYPT = 48 };

gcry_error_t gcry_md_open(gcry_md_hd_t *h, int algo, unsigned int flags);
void gcry_md_write(gcry_md_hd_t h, const void *buf, size_t len);
unsigned char *gcry_md_read(gcry_md_hd_t h, int algo);
gcry_error_t gcry_md_extract(gcry_md_hd_t h, int algo, void *buf, size_t len);
void gcry_md_reset(gcry_md_hd_t h);
void gcry_md_close(gcry_md_hd_t h);
void gcry_create_nonce(void *buf, size_t len);
gcry_error_t gcry_cipher_open(gcry_cipher_hd_t *h, int algo, int mode, unsigned int flags);
gcry_error_t gcry_cipher_setkey(gcry_cipher_hd_t h, const void *key, size_t len);
gcry_error_t gcry_cipher_setiv(gcry_cipher_hd_t h, const void *iv, size_t len);
gcry_error_t gcry_cipher_encrypt(gcry_cipher_hd_t h, void *out, size_t outsize, const void *in, size_t inlen);
gcry_error_t gcry_cipher_decrypt(gcry_cipher_hd_t h, void *out, size_t outsize, const void *in, size_t inlen);
void gcry_cipher_close(gcry_cipher_hd_t h);
gcry_error_t gcry_kdf_derive(const void *pass, size_t passlen, int algo, int subalgo,
                             const void *salt, size_t saltlen, unsigned long iterations,
                             size_t keysize, void *keybuffer);
#ifdef __cplusplus
}
#endif
#endif
