/* c_abi_smoke.c -- libsa_cuda.so used from plain C (no Python, no torch): the drop-in boundary is a
 * C ABI.  Without a CUDA device sa_create must fail with SA_ERR_CUDA and a message (there is no CPU
 * fallback) -> prints "nogpu" and exits 0.  With a device: builds the index of "banana\nbandana",
 * checks the array and two query ranges against the reference's known answers (the same
 * vectors as tests/test_oracle_vs_ref.py::test_known_answer_banana) -> prints "ok".
 *   gcc -Iinclude tests/c_abi_smoke.c -Lsuffixarray_b200 -lsa_cuda -Wl,-rpath,$PWD/suffixarray_b200 -o /tmp/c_abi_smoke */
#include <stdio.h>
#include <string.h>

#include "sa_cuda.h"

int main(void)
{
    if (sa_abi_version() != SA_CUDA_ABI_VERSION) { printf("abi version mismatch\n"); return 2; }
    sa_index* idx = NULL;
    int rc = sa_create(&idx, 32, 0);
    if (rc == SA_ERR_CUDA) { printf("nogpu: %s\n", sa_last_error()); return idx == NULL ? 0 : 3; }
    if (rc != SA_OK) { printf("sa_create: %d %s\n", rc, sa_last_error()); return 4; }
    const char* text = "banana\nbandana";
    const uint64_t n = 14;
    rc = sa_build_text(idx, (const uint8_t*)text, n, 0);
    if (rc != SA_OK) { printf("sa_build_text: %d %s\n", rc, sa_last_error()); return 5; }
    static const uint32_t want[14] = {6, 13, 5, 11, 3, 1, 8, 0, 7, 10, 12, 4, 2, 9};
    uint32_t sa[14];
    if (sa_size(idx) != n || sa_copy_sa_u32(idx, sa, n) != SA_OK || memcmp(sa, want, sizeof want) != 0) {
        printf("suffix array differs\n");
        return 6;
    }
    const uint8_t q[] = "anaban";
    const uint64_t qoff[3] = {0, 3, 6};
    uint32_t start[2], end[2];
    uint64_t hit_off[3];
    rc = sa_query_ranges(idx, q, qoff, 2, 1000, start, end, hit_off);
    if (rc != SA_OK || start[0] != 3 || end[0] != 5 || start[1] != 7 || end[1] != 8 || hit_off[2] != 5) {
        printf("query ranges differ: %d [%u,%u] [%u,%u] hits %llu\n", rc, start[0], end[0], start[1], end[1],
               (unsigned long long)hit_off[2]);
        return 7;
    }
    uint32_t pos[5], doc[5];
    rc = sa_query_hits(idx, pos, doc, 5);
    /* "ana": ranks 3..5 = positions 11, 3, 1 -> records 1, 0, 0; "ban": ranks 7..8 = positions 0, 7 -> records 0, 1 */
    static const uint32_t want_pos[5] = {11, 3, 1, 0, 7}, want_doc[5] = {1, 0, 0, 0, 1};
    if (rc != SA_OK || memcmp(pos, want_pos, sizeof pos) != 0 || memcmp(doc, want_doc, sizeof doc) != 0) {
        printf("hits differ\n");
        return 8;
    }
    /* ABI 2: a per-call result object, de-duplicated hits, records cut in the library */
    sa_result* res = NULL;
    if (sa_result_create(idx, &res) != SA_OK) { printf("sa_result_create: %s\n", sa_last_error()); return 9; }
    rc = sa_query_ranges_r(idx, res, q, qoff, 2, 1000, 0, start, end, hit_off);
    uint64_t doff[3];
    if (rc == SA_OK) rc = sa_query_hits_r(idx, res, SA_QUERY_DEDUPE, pos, doc, 5, doff);
    /* "ana" -> records 1, 0 (the second hit in record 0 is dropped); "ban" -> records 0, 1 */
    static const uint32_t want_ddoc[4] = {1, 0, 0, 1};
    if (rc != SA_OK || doff[1] != 2 || doff[2] != 4 || memcmp(doc, want_ddoc, sizeof want_ddoc) != 0) {
        printf("de-duplicated hits differ: %d %s\n", rc, sa_last_error());
        return 10;
    }
    uint64_t nrec = 0, nbytes = 0, rec_off[6];
    uint8_t recs[64];
    rc = sa_query_records_size(idx, res, 0, &nrec, &nbytes);
    if (rc == SA_OK && nrec == 5 && nbytes <= sizeof recs) rc = sa_query_records(idx, res, recs, sizeof recs, rec_off, 6, NULL, NULL);
    /* records of the five hits: bandana, banana, banana, banana, bandana */
    if (rc != SA_OK || nrec != 5 || nbytes != 7 + 6 + 6 + 6 + 7 || memcmp(recs, "bandanabananabananabananabandana", 32) != 0) {
        printf("records differ: %d n=%llu bytes=%llu %s\n", rc, (unsigned long long)nrec, (unsigned long long)nbytes, sa_last_error());
        return 11;
    }
    sa_result_destroy(res);
    sa_destroy(idx);
    printf("ok\n");
    return 0;
}
