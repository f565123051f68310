/* oracle/fitch_oracle.h — TEST INFRASTRUCTURE ONLY (see fitch_oracle.c). */
#ifndef LVB_B200_FITCH_ORACLE_H
#define LVB_B200_FITCH_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

long oracle_words_per_row(long m);
int oracle_dna_to_binary(long n, long m, const char *const *rows, uint64_t *out);
long oracle_variable_columns(long n, long m, const char *const *rows, unsigned char *keep);
long oracle_min_tree_length(long n, long m, const char *const *rows);
int oracle_fitch_word(uint64_t x, uint64_t y, uint64_t *out);
long oracle_getplen(long n, long nwords, const long *left, const long *right, long root,
                    uint64_t *rows, long *changes);
long oracle_getplen_sliced(long n, long nwords, const long *left, const long *right, long root,
                           uint64_t *rows, long *changes, int nslices);
void oracle_dirty_all(long n, long nwords, uint64_t *rows);

#ifdef __cplusplus
}
#endif
#endif
