/* oracle/orc_types.h -- CPU ORACLE, TEST INFRASTRUCTURE ONLY: the read model shared by edge_oracle.c and quasi_oracle.c. */
#ifndef ORC_TYPES_H_
#define ORC_TYPES_H_
#include <stdint.h>

typedef struct {
    int32_t start, end;      /* 1-based inclusive, as AlignmentRecord::start/end (src/AlignmentRecord.cpp:190-191) */
    int32_t ncig, nq;        /* unrolled CIGAR length, query length */
    const char* cig;         /* unrolled CIGAR chars: M I D S H */
    const char* seq;         /* bases, already normalised to ACGTN (src/ShortDnaSequence.cpp:37-50) */
    const char* qual;        /* ASCII Phred+33 */
} orc_seg;

typedef struct {
    int32_t single_end;      /* AlignmentRecord::single_end */
    int32_t is_clique;       /* name contains "Clique" (src/NewEdgeCalculator.cpp:1111) */
    int32_t ncov;            /* cov_pos.size() (:1127) */
    int32_t cov_first, cov_last; /* cov_pos.front().ref / back().ref (:1252) */
    orc_seg s[2];
} orc_read;

#endif
