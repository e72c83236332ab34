/*
 * ref_shim.c -- thin C wrapper that links the UNMODIFIED reference sources
 * (/root/reference/suffix_array/engine.c, libsais.c) into oracle/_ref/libsa_ref.so so that
 * tests and bench.py can call them through ctypes.  TEST INFRASTRUCTURE ONLY: nothing under
 * suffixarray_b200/ loads this.  No reference source is copied; oracle/Makefile compiles the
 * files where they lie.  All functions here only marshal arguments and silence the
 * reference's printf()s; the arithmetic is the reference's own.
 */
#define _GNU_SOURCE
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <unistd.h>
#include <fcntl.h>
#include <omp.h>
#include <ctype.h>
#include "engine.h"
#include "libsais.h"
#include "libsais64.h"

static int g_saved_fd = -1;
static void hush(void)
{
    fflush(stdout);
    g_saved_fd = dup(1);
    int dn = open("/dev/null", O_WRONLY);
    if (dn >= 0) { dup2(dn, 1); close(dn); }
}
static void unhush(void)
{
    fflush(stdout);
    if (g_saved_fd >= 0) { dup2(g_saved_fd, 1); close(g_saved_fd); g_saved_fd = -1; }
}

/* construct_truncated_suffix_array (engine.c:837) on caller-owned buffers.
 * text_padded must have >= 128 readable zero bytes after n. */
void ref_build(const char* text_padded, uint32_t n, uint32_t max_suffix_length, uint32_t* sa_out)
{
    SuffixArray_struct s;
    memset(&s, 0, sizeof s);
    s.suffix_array = sa_out;
    s.max_suffix_length = max_suffix_length;
    s.n = n;
    s.global_byte_end_idx = n;
    hush();
    construct_truncated_suffix_array(text_padded, &s);
    unhush();
}

/* get_substring_positions (engine.c:869).  Returns 1 and leaves outputs untouched when the
 * call could take the reference's `last = mid - 1` wrap at mid == 0 (out-of-bounds read):
 * that needs strncmp(text + SA[0], q, c) >= 0, so such queries are simply not submitted. */
int ref_query(const char* text_padded, const uint32_t* sa, uint32_t n, uint32_t max_suffix_length,
              const char* q_cstr, uint32_t* start, uint32_t* end)
{
    SuffixArray_struct s;
    memset(&s, 0, sizeof s);
    s.suffix_array = (uint32_t*)sa;
    s.max_suffix_length = max_suffix_length;
    s.n = n;
    if (n == 0) return 1;
    size_t m = strlen(q_cstr);
    size_t c = m < max_suffix_length ? m : max_suffix_length;
    if (strncmp(text_padded + sa[0], q_cstr, c) >= 0) return 1;
    pair_u32 r = get_substring_positions(text_padded, &s, q_cstr);
    *start = r.first; *end = r.second;
    return 0;
}

/* Batch of NUL-terminated queries at qbytes + qoff[i]; skipped[i] = 1 where ref_query refused.
 * nthreads <= 0: all cores (an omp parallel-for over the batch, SURVEY 8(d) baseline A). */
void ref_query_batch(const char* text_padded, const uint32_t* sa, uint32_t n, uint32_t max_suffix_length,
                     const char* qbytes, const uint64_t* qoff, uint64_t nq,
                     uint32_t* start, uint32_t* end, uint8_t* skipped, int nthreads)
{
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(static) num_threads(nthreads)
    for (long i = 0; i < (long)nq; ++i) {
        skipped[i] = (uint8_t)ref_query(text_padded, sa, n, max_suffix_length, qbytes + qoff[i],
                                        &start[i], &end[i]);
    }
}

/* get_matching_records (engine.c:1168-1214), the reference's in-memory "query -> records" path, for
 * a batch under an omp parallel-for: bisection + for each of the first min(k, count) hits the
 * newline scans, malloc and memcpy of the record.  The records are freed at once; the return value
 * is the total number of records produced (hits[i] = per query).  The caller must provide a '\n'
 * at text_padded[-1] (the backward scan of a hit in record 0 otherwise leaves the buffer,
 * engine.c:1190) and a '\n' somewhere after the >= 64 zero bytes that follow the text (the forward
 * scan of the last record, engine.c:1194-1205), plus max-record-length readable bytes after that
 * (the memcpy at engine.c:1208 copies from AFTER the record).  Queries the shim refuses
 * (out-of-bounds wrap, see ref_query) produce no records. */
uint64_t ref_query_records_batch(const char* text_padded, const uint32_t* sa, uint32_t n,
                                 uint32_t max_suffix_length, const char* qbytes, const uint64_t* qoff,
                                 uint64_t nq, uint32_t k, uint32_t* hits, int nthreads)
{
    if (nthreads <= 0) nthreads = omp_get_max_threads();
    uint64_t total = 0;
#pragma omp parallel num_threads(nthreads) reduction(+ : total)
    {
        char** recs = (char**)malloc((size_t)k * sizeof(char*));
        SuffixArray_struct s;
        memset(&s, 0, sizeof s);
        s.suffix_array = (uint32_t*)sa;
        s.max_suffix_length = max_suffix_length;
        s.n = n;
#pragma omp for schedule(dynamic, 64)
        for (long i = 0; i < (long)nq; ++i) {
            const char* q = qbytes + qoff[i];
            size_t m = strlen(q);
            size_t c = m < max_suffix_length ? m : max_suffix_length;
            hits[i] = 0;
            if (n == 0 || strncmp(text_padded + sa[0], q, c) >= 0) continue;
            uint32_t got = get_matching_records(text_padded, &s, q, k, recs);
            for (uint32_t j = 0; j < got; ++j) free(recs[j]);
            hits[i] = got;
            total += got;
        }
        free(recs);
    }
    return total;
}

/* CSV path end to end: construct_truncated_suffix_array_from_csv_partitioned_mmap_full
 * (engine.c:1454).  Single-partition files only (< 2 GiB).  Returns n, or 0 when the
 * column was not found; copies the SA (file offsets) and the per-rank quoted bits
 * (one byte per rank) into caller buffers of capacity cap. */
uint32_t ref_csv_build(const char* path, const char* column, uint32_t max_suffix_length,
                       uint32_t* sa_out, uint8_t* quoted_out, uint64_t cap)
{
    SuffixArrayIndex* idx = NULL;
    char* cols[1];
    cols[0] = (char*)column;
    hush();
    construct_truncated_suffix_array_from_csv_partitioned_mmap_full(&idx, path, max_suffix_length, cols, 1);
    unhush();
    SuffixArray_struct* s = &idx->suffix_arrays[0];
    uint32_t n = s->n;
    if (n <= cap) {
        memcpy(sa_out, s->suffix_array, (size_t)n * sizeof(uint32_t));
        for (uint32_t i = 0; i < n; ++i) quoted_out[i] = get_buffer_bit(s->is_quoted_bitflag, i);
    }
    free(s->suffix_array);
    return n;
}

/* ---- the reference's CSV query path on an index it built itself ------------------------------
 * ref_csv_open  = construct_truncated_suffix_array_from_csv_partitioned_mmap_full (engine.c:1454)
 * ref_csv_query = get_substring_positions_file (engine.c:920-999) on partition 0
 * ref_csv_records = get_matching_records_file (engine.c:1326-1390)
 * Single-partition files only (< 2 GiB).  Two fences, both around undefined behaviour of the
 * reference (out-of-bounds reads, use of a closed FILE): ref_csv_query returns 1 without calling it
 * when the probe at rank 0 would compare >= 0 (`last = mid - 1` wraps at mid == 0, :951), and
 * ref_csv_records only calls get_matching_records_file for queries that match and whose hits all
 * lie behind the header line (the backward newline seek of a hit in row 0 returns UINT32_MAX, :1283). */
void* ref_csv_open(const char* path, const char* column, uint32_t max_suffix_length)
{
    SuffixArrayIndex* idx = NULL;
    char* cols[1];
    cols[0] = (char*)column;
    hush();
    construct_truncated_suffix_array_from_csv_partitioned_mmap_full(&idx, path, max_suffix_length, cols, 1);
    unhush();
    return idx;
}

uint32_t ref_csv_n(void* h) { return ((SuffixArrayIndex*)h)->suffix_arrays[0].n; }

/* files >= 2 GiB: one SuffixArray_struct per 2 GiB partition of the FILE (engine.c:1437-1438, :1469-1481) */
uint32_t ref_csv_partitions(void* h) { return ((SuffixArrayIndex*)h)->num_partitions; }
uint32_t ref_csv_part_n(void* h, uint32_t p) { return ((SuffixArrayIndex*)h)->suffix_arrays[p].n; }
void ref_csv_part_bounds(void* h, uint32_t p, uint64_t* start, uint64_t* end)
{
    *start = ((SuffixArrayIndex*)h)->suffix_arrays[p].global_byte_start_idx;
    *end = ((SuffixArrayIndex*)h)->suffix_arrays[p].global_byte_end_idx;
}
void ref_csv_part_copy(void* h, uint32_t p, uint32_t* sa_out, uint8_t* quoted_out)
{
    SuffixArray_struct* s = &((SuffixArrayIndex*)h)->suffix_arrays[p];
    memcpy(sa_out, s->suffix_array, (size_t)s->n * sizeof(uint32_t));
    if (quoted_out)
        for (uint32_t i = 0; i < s->n; ++i) quoted_out[i] = get_buffer_bit(s->is_quoted_bitflag, i);
}

void ref_csv_copy(void* h, uint32_t* sa_out, uint8_t* quoted_out)
{
    SuffixArray_struct* s = &((SuffixArrayIndex*)h)->suffix_arrays[0];
    memcpy(sa_out, s->suffix_array, (size_t)s->n * sizeof(uint32_t));
    for (uint32_t i = 0; i < s->n; ++i) quoted_out[i] = get_buffer_bit(s->is_quoted_bitflag, i);
}

void ref_csv_close(void* h)
{
    SuffixArrayIndex* idx = (SuffixArrayIndex*)h;
    free(idx->suffix_arrays[0].suffix_array);
}

static int rank0_wraps(FILE* f, const SuffixArray_struct* s, const char* q)
{
    size_t m = strlen(q);
    size_t c = m < s->max_suffix_length ? m : s->max_suffix_length;
    char* line = (char*)calloc(c + 8, 1);
    fseek(f, (long)(s->global_byte_start_idx + s->suffix_array[0]), SEEK_SET);
    fread(line, 1, c, f);
    for (size_t i = 0; i < c; ++i) line[i] = (char)tolower(line[i]);
    int r = strncmp(line, q, c) >= 0;
    free(line);
    return r;
}

int ref_csv_query(void* h, const char* path, const char* q, uint32_t* start, uint32_t* end)
{
    SuffixArray_struct* s = &((SuffixArrayIndex*)h)->suffix_arrays[0];
    FILE* f = fopen(path, "r");
    if (!f || s->n == 0) { if (f) fclose(f); return 1; }
    if (rank0_wraps(f, s, q)) { fclose(f); return 1; }
    pair_u32 r = get_substring_positions_file(f, s, q);
    fclose(f);
    *start = r.first; *end = r.second;
    return 0;
}

/* Batch of NUL-terminated queries under an omp parallel-for (each thread its own FILE*). */
void ref_csv_query_batch(void* h, const char* path, const char* qbytes, const uint64_t* qoff, uint64_t nq,
                         uint32_t* start, uint32_t* end, uint8_t* skipped, int nthreads)
{
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 16) num_threads(nthreads)
    for (long i = 0; i < (long)nq; ++i)
        skipped[i] = (uint8_t)ref_csv_query(h, path, qbytes + qoff[i], &start[i], &end[i]);
}

/* -> number of records (>= 0), -1 when the query was fenced.  Records are copied into out
 * (rec_off[j] .. rec_off[j + 1]) and freed. */
int64_t ref_csv_records(void* h, const char* path, const char* q, uint32_t k, char* out, uint64_t cap,
                        uint64_t* rec_off)
{
    SuffixArray_struct* s = &((SuffixArrayIndex*)h)->suffix_arrays[0];
    uint32_t start = UINT32_MAX, end = UINT32_MAX;
    rec_off[0] = 0;
    if (ref_csv_query(h, path, q, &start, &end)) return -1;
    if (start == UINT32_MAX || end == UINT32_MAX || start == end + 1) return 0;      /* the reference fclose()s and goes on, :1346-1356 */
    uint32_t take = end - start + 1 < k ? end - start + 1 : k;
    FILE* f = fopen(path, "r");
    if (!f) return -1;
    long header_end = 0;
    int ch;
    while ((ch = fgetc(f)) != EOF && ch != '\n') ++header_end;
    fclose(f);
    for (uint32_t i = 0; i < take; ++i)
        if ((long)s->suffix_array[start + i] <= header_end) return -1;
    char** recs = (char**)calloc(k ? k : 1, sizeof(char*));
    uint32_t got = 0;
    hush();
    get_matching_records_file(path, s, q, k, recs, &got);
    unhush();
    uint64_t w = 0;
    for (uint32_t j = 0; j < got; ++j) {
        size_t len = strlen(recs[j]);
        if (w + len <= cap) memcpy(out + w, recs[j], len);
        w += len;
        rec_off[j + 1] = w;
        free(recs[j]);
    }
    free(recs);
    return (w <= cap) ? (int64_t)got : -1;
}

/* libsais (vendored 2.8.4), the "baseline B" timing arm -- a FULL suffix array, never a
 * parity oracle.  threads <= 1: libsais(); else libsais_omp(). */
int ref_libsais(const uint8_t* text, int32_t* sa, int32_t n, int threads)
{
#ifdef LIBSAIS_OPENMP
    if (threads != 1) return libsais_omp(text, sa, n, 0, NULL, threads < 0 ? 0 : threads);
#endif
    (void)threads;
    return libsais(text, sa, n, 0, NULL);
}

/* libsais64 (vendored, /root/reference/suffix_array/libsais64.h:61,:86): the int64 arm BASELINE.json names
 * for the 4 G-char config; never called by the reference's live path.  Timing baseline only. */
int64_t ref_libsais64(const uint8_t* text, int64_t* sa, int64_t n, int threads)
{
#ifdef LIBSAIS_OPENMP
    if (threads != 1) return libsais64_omp(text, sa, n, 0, NULL, threads < 0 ? 0 : threads);
#endif
    (void)threads;
    return libsais64(text, sa, n, 0, NULL);
}

int ref_max_threads(void) { return omp_get_max_threads(); }

/* torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the timed CPU arm sets the
 * thread count explicitly (engine.c's own `#pragma omp parallel for`s take the default). */
void ref_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
