/* host_internal.h -- shared by the host C files, not part of the ABI */
#ifndef FEM_HOST_INTERNAL_H_
#define FEM_HOST_INTERNAL_H_
#include <stddef.h>
#include <stdint.h>

#include "femhost.h"

int femhost_set_error(int code, const char *fmt, ...);
int femhost_records_reserve(femhost_records *r, size_t n_records, size_t n_bases, size_t n_names);
int femhost_str_reserve(femhost_str *s, size_t extra);
double femhost_now(void);
int femhost_parse_fastq4(const unsigned char *p, const unsigned char *end, femhost_records *out);
long femhost_fastq4_sync(const unsigned char *buf, size_t n, int at_file_start, int at_file_end);

/* src/utils.h:72: A/a 0, C/c 1, G/g 2, T/t 3, everything else 4.  The table stores code+1 so that
 * its zero default means "ambiguous"; read it through FEM_CODE(). */
extern const uint8_t femhost_base_code[256];
#define FEM_CODE(c) ((uint8_t)(femhost_base_code[(uint8_t)(c)] ? femhost_base_code[(uint8_t)(c)] - 1 : 4))
#endif
