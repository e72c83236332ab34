/*
 * sa_oracle.c -- CPU restatement of the reference's hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is the parity checker for libsa_cuda.so.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it; the product
 * (suffixarray_b200/) never does and has no CPU fallback.
 *
 * It restates, in plain scalar C (no SSE, no recursion), what the reference computes:
 *   - ORDER-E, the order produced by construct_truncated_suffix_array
 *       /root/reference/suffix_array/engine.c:837-866 (driver), :696-812 (recursive_bucket_sort),
 *       :664-677 (strncmp_128, the 17-byte window compare of the <=32 insertion sort);
 *   - the two fixed-probe bisections of get_substring_positions, engine.c:869-918;
 *   - hit expansion: first min(k, count) ranks in SA order, engine.c:1185-1187, with the
 *     record of a hit identified by the number of '\n' before it (engine.c:1189-1191 scans
 *     back to the previous '\n'; the count is the same record, without the doc-0 underflow);
 *   - the CSV column scan of construct_truncated_suffix_array_from_csv_partitioned_mmap,
 *     engine.c:522-605, and the post-build remap, engine.c:632-651.
 *
 * Parity status: PINNED.  The reference ships no golden vectors (tests/test.py has no
 * asserts), so this restatement is pinned against the reference itself, compiled unmodified
 * from /root/reference into oracle/_ref/libsa_ref.so (oracle/Makefile), on the seeded
 * matrix in tests/test_oracle_vs_ref.py, and against the vectors that run produced and
 * committed under tests/golden/ (tests/golden/make_golden.py).
 *
 * Defined behaviour where the reference has none:
 *   - the text is followed by zero bytes (the reference reads past the end of an
 *     unterminated buffer, engine.c:738,752 and :664-666; zero is what calloc/mmap gives);
 *   - `last = mid - 1` at mid == 0 (engine.c:891,907) wraps to UINT32_MAX in the reference and
 *     reads out of bounds; here the loop ends (signed arithmetic).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define SAO_PAD 512u          /* zero bytes kept after the text */
#define SAO_SMALL 32u         /* insertion-sort cut-off, engine.c:710 */

/* ---- 17-byte window compare: restates strncmp_128, engine.c:664-677 ------------------------
 * pcmpistri EQUAL_EACH|NEGATIVE_POLARITY with implicit lengths: index of the first byte in
 * 0..15 where the windows differ, a byte past a window's first NUL being "invalid"
 * (invalid vs invalid = equal, invalid vs valid = differ); 16 if none.  The result is the
 * raw byte difference AT that index -- index 16 is the 17th byte, outside the window --
 * gated by idx < max_depth (the build's L' = min(L, n), NOT L' - depth). */
static int w17_compare(const uint8_t* a, const uint8_t* b, uint32_t lprime)
{
    uint32_t idx = 16;
    for (uint32_t i = 0; i < 16; ++i) {
        if (a[i] != b[i]) { idx = i; break; }
        if (a[i] == 0) break;            /* both windows end here: every later lane is equal */
    }
    if (idx >= lprime) return 0;
    return (int)a[idx] - (int)b[idx];
}

typedef struct { uint32_t lo, hi, depth; } sao_frame;

typedef struct {
    sao_frame* v;
    size_t n, cap;
} sao_stack;

static int stack_push(sao_stack* s, uint32_t lo, uint32_t hi, uint32_t depth)
{
    if (s->n == s->cap) {
        size_t nc = s->cap ? s->cap * 2 : 1024;
        sao_frame* nv = (sao_frame*)realloc(s->v, nc * sizeof(sao_frame));
        if (!nv) return -1;
        s->v = nv; s->cap = nc;
    }
    s->v[s->n].lo = lo; s->v[s->n].hi = hi; s->v[s->n].depth = depth;
    ++s->n;
    return 0;
}

/* One bucket [lo,hi) at `depth`: restates one activation of recursive_bucket_sort
 * (engine.c:696-812) and pushes the sub-buckets it would recurse into. */
static int refine_bucket(const uint8_t* t, uint32_t* sa, uint32_t* tmp, sao_frame f,
                         uint32_t lprime, sao_stack* out)
{
    const uint32_t size = f.hi - f.lo;
    uint32_t* b = sa + f.lo;
    if (f.depth == lprime) return 0;                       /* engine.c:706 */
    if (size <= SAO_SMALL) {                               /* engine.c:710-729 */
        for (uint32_t i = 1; i < size; ++i) {
            uint32_t j = i;
            while (j > 0 && w17_compare(t + b[j] + f.depth, t + b[j - 1] + f.depth, lprime) < 0) {
                uint32_t x = b[j]; b[j] = b[j - 1]; b[j - 1] = x;
                --j;
            }
        }
        return 0;
    }
    /* stable counting sort on byte `depth` (engine.c:734-758) */
    uint32_t count[256];
    uint32_t start[257];
    memset(count, 0, sizeof count);
    for (uint32_t i = 0; i < size; ++i) ++count[t[(uint64_t)b[i] + f.depth]];
    start[0] = 0;
    for (int c = 0; c < 256; ++c) start[c + 1] = start[c] + count[c];
    uint32_t cursor[256];
    memcpy(cursor, start, sizeof cursor);
    uint32_t* o = tmp + f.lo;
    for (uint32_t i = 0; i < size; ++i) o[cursor[t[(uint64_t)b[i] + f.depth]]++] = b[i];
    memcpy(b, o, (size_t)size * sizeof(uint32_t));
    /* children: bytes 0..254 except '\n'; the '\n' and 0xFF buckets are never refined
     * (engine.c:771-773, :795-796: loop bound 255 and the '\n' skip) */
    for (int c = 0; c < 255; ++c) {
        if (c == '\n') continue;
        if (count[c] <= 1) continue;
        if (stack_push(out, f.lo + start[c], f.lo + start[c + 1], f.depth + 1)) return -1;
    }
    return 0;
}

static uint8_t* padded_copy(const uint8_t* text, uint64_t n)
{
    uint8_t* t = (uint8_t*)calloc(n + SAO_PAD, 1);
    if (t && n) memcpy(t, text, n);
    return t;
}

/* ORDER-E.  text: n bytes (no padding needed; a zero-padded copy is made).  sa: n entries.
 * Restates construct_truncated_suffix_array, engine.c:837-866.  Returns 0, or -1 on OOM,
 * -2 on n >= 2^32-1 or L == 0 handled as: L == 0 gives the identity (depth 0 == max depth). */
int sao_build_tail(const uint8_t* text, uint64_t n, uint64_t tail, uint32_t max_suffix_length, uint32_t* sa);
int sao_build(const uint8_t* text, uint64_t n, uint32_t max_suffix_length, uint32_t* sa)
{
    return sao_build_tail(text, n, 0, max_suffix_length, sa);
}

/* One partition of a longer text: the reference calls construct_truncated_suffix_array on
 * `text + idx * 2 GiB` of one contiguous buffer (suffix_array.pyx:176-180), so the compares of the
 * last suffixes read on into the next partition.  text holds n + tail bytes, n are indexed. */
int sao_build_tail(const uint8_t* text, uint64_t n, uint64_t tail, uint32_t max_suffix_length, uint32_t* sa)
{
    if (n >= 0xFFFFFFFFull) return -2;
    if (n == 0) return 0;
    uint8_t* t = padded_copy(text, n + tail);
    uint32_t* tmp = (uint32_t*)malloc((size_t)n * sizeof(uint32_t));
    if (!t || !tmp) { free(t); free(tmp); return -1; }
    const uint32_t lprime = max_suffix_length < n ? max_suffix_length : (uint32_t)n;   /* :841 */
    for (uint64_t i = 0; i < n; ++i) sa[i] = (uint32_t)i;                                /* :845-848 */

    int rc = 0;
    /* depth 0 runs alone, then its sub-buckets are independent (the reference runs them
     * under omp parallel for schedule(dynamic), engine.c:768-791); each thread drains its
     * own stack so the result does not depend on the thread count. */
    sao_stack top = {0, 0, 0};
    sao_frame root = {0, (uint32_t)n, 0};
    rc = refine_bucket(t, sa, tmp, root, lprime, &top);
    if (rc == 0) {
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic)
#endif
        for (long k = 0; k < (long)top.n; ++k) {
            sao_stack st = {0, 0, 0};
            int bad = stack_push(&st, top.v[k].lo, top.v[k].hi, top.v[k].depth);
            while (!bad && st.n) {
                sao_frame f = st.v[--st.n];
                bad = refine_bucket(t, sa, tmp, f, lprime, &st);
            }
            free(st.v);
            if (bad) {
#ifdef _OPENMP
#pragma omp atomic write
#endif
                rc = -1;
            }
        }
    }
    free(top.v);
    free(tmp);
    free(t);
    return rc;
}

/* strncmp(text + pos, q, c) where q has no NUL in its first c bytes (c <= strlen(q)):
 * byte-wise unsigned compare, which therefore also stops at a NUL in the text. */
static int cmp_window(const uint8_t* t, uint32_t pos, const uint8_t* q, uint32_t c)
{
    const uint8_t* a = t + pos;
    for (uint32_t i = 0; i < c; ++i) {
        if (a[i] != q[i]) return (int)a[i] - (int)q[i];
    }
    return 0;
}

/* Restates get_substring_positions, engine.c:869-918.  `t` must be zero-padded. */
static void query_padded(const uint8_t* t, uint64_t n, const uint32_t* sa, uint32_t max_suffix_length,
                         const uint8_t* q, uint32_t qlen, uint32_t* out_start, uint32_t* out_end)
{
    uint32_t m = 0;
    while (m < qlen && q[m] != 0) ++m;                           /* strlen, :874 */
    const uint32_t c = m < max_suffix_length ? m : max_suffix_length;   /* :881 (L, not L') */
    uint32_t start = UINT32_MAX, end = UINT32_MAX;
    if (n == 0) { *out_start = start; *out_end = end; return; }
    int64_t first = 0, last = (int64_t)n - 1;
    while (first <= last) {                                      /* :884-894 */
        int64_t mid = (first + last) / 2;
        if (cmp_window(t, sa[mid], q, c) < 0) first = mid + 1;
        else { last = mid - 1; start = (uint32_t)mid; }
    }
    if (start == UINT32_MAX) { *out_start = start; *out_end = end; return; }   /* :896-899 */
    first = 0; last = (int64_t)n - 1;
    while (first <= last) {                                      /* :902-914 */
        int64_t mid = (first + last) / 2;
        if (cmp_window(t, sa[mid], q, c) > 0) last = mid - 1;
        else { first = mid + 1; end = (uint32_t)mid; }
    }
    *out_start = start; *out_end = end;
}

/* Batched ranges.  queries are qbytes[qoff[i] .. qoff[i+1]).  nthreads <= 0: all cores. */
int sao_query_batch(const uint8_t* text, uint64_t n, const uint32_t* sa, uint32_t max_suffix_length,
                    const uint8_t* qbytes, const uint64_t* qoff, uint64_t nq,
                    uint32_t* start, uint32_t* end, int nthreads)
{
    uint8_t* t = padded_copy(text, n);
    if (!t) return -1;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(static) num_threads(nthreads)
#endif
    for (long i = 0; i < (long)nq; ++i) {
        query_padded(t, n, sa, max_suffix_length, qbytes + qoff[i], (uint32_t)(qoff[i + 1] - qoff[i]),
                     &start[i], &end[i]);
    }
    (void)nthreads;
    free(t);
    return 0;
}

/* Partition variant: text holds n + tail bytes (see sao_build_tail); the bisection runs over the n
 * entries of sa and its compares read the look-ahead bytes. */
int sao_query_batch_tail(const uint8_t* text, uint64_t n, uint64_t tail, const uint32_t* sa, uint32_t max_suffix_length,
                         const uint8_t* qbytes, const uint64_t* qoff, uint64_t nq, uint32_t* start, uint32_t* end)
{
    uint8_t* t = padded_copy(text, n + tail);
    if (!t) return -1;
    for (uint64_t i = 0; i < nq; ++i)
        query_padded(t, n, sa, max_suffix_length, qbytes + qoff[i], (uint32_t)(qoff[i + 1] - qoff[i]), &start[i], &end[i]);
    free(t);
    return 0;
}

/* Same, but the caller guarantees `text` already has >= max_suffix_length zero bytes after n
 * (used by the timed CPU baseline so the copy is not inside the timed region). */
int sao_query_batch_padded(const uint8_t* text_padded, uint64_t n, const uint32_t* sa,
                           uint32_t max_suffix_length, const uint8_t* qbytes, const uint64_t* qoff,
                           uint64_t nq, uint32_t* start, uint32_t* end, int nthreads)
{
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(static) num_threads(nthreads)
#endif
    for (long i = 0; i < (long)nq; ++i) {
        query_padded(text_padded, n, sa, max_suffix_length, qbytes + qoff[i],
                     (uint32_t)(qoff[i + 1] - qoff[i]), &start[i], &end[i]);
    }
    (void)nthreads;
    return 0;
}

/* Number of hits the reference materialises for a range: engine.c:1181-1185
 * (first == UINT32_MAX -> 0, else min(k, second - first + 1) in uint32 arithmetic). */
static uint32_t take_count(uint32_t start, uint32_t end, uint32_t k)
{
    if (start == UINT32_MAX) return 0;
    uint32_t cnt = end - start + 1u;
    return cnt < k ? cnt : k;
}

/* hit_off[nq+1] = exclusive prefix sum of take counts. */
void sao_hit_offsets(const uint32_t* start, const uint32_t* end, uint64_t nq, uint32_t k, uint64_t* hit_off)
{
    uint64_t acc = 0;
    for (uint64_t i = 0; i < nq; ++i) { hit_off[i] = acc; acc += take_count(start[i], end[i], k); }
    hit_off[nq] = acc;
}

/* Hit expansion: for query i, ranks start[i] .. start[i]+take-1 in rank order, duplicates kept
 * (engine.c:1187).  hit_pos = SA[rank]; hit_doc = number of '\n' in text[0 .. pos). */
int sao_hits(const uint8_t* text, uint64_t n, const uint32_t* sa, const uint32_t* start,
             const uint32_t* end, uint64_t nq, uint32_t k, const uint64_t* hit_off,
             uint32_t* hit_pos, uint32_t* hit_doc)
{
    /* newline positions, ascending */
    uint64_t nl = 0;
    for (uint64_t i = 0; i < n; ++i) nl += (text[i] == '\n');
    uint32_t* nlpos = (uint32_t*)malloc((size_t)(nl ? nl : 1) * sizeof(uint32_t));
    if (!nlpos) return -1;
    uint64_t w = 0;
    for (uint64_t i = 0; i < n; ++i) if (text[i] == '\n') nlpos[w++] = (uint32_t)i;
    for (uint64_t i = 0; i < nq; ++i) {
        uint32_t take = take_count(start[i], end[i], k);
        for (uint32_t j = 0; j < take; ++j) {
            uint32_t p = sa[(uint64_t)start[i] + j];
            /* doc = #newlines strictly before p = lower_bound(nlpos, p) */
            uint64_t lo = 0, hi = nl;
            while (lo < hi) { uint64_t mid = (lo + hi) / 2; if (nlpos[mid] < p) lo = mid + 1; else hi = mid; }
            hit_pos[hit_off[i] + j] = p;
            hit_doc[hit_off[i] + j] = (uint32_t)lo;
        }
    }
    free(nlpos);
    return 0;
}

static uint8_t lower_ascii(uint8_t c) { return (c >= 'A' && c <= 'Z') ? (uint8_t)(c + 32) : c; }

/* ASCII lower-casing of the joined text: restates lowercase_string,
 * /root/reference/suffix_array/suffix_array.pyx:103-107 (65..90 -> +32, nothing else). */
void sao_lowercase(uint8_t* s, uint64_t n)
{
    for (uint64_t i = 0; i < n; ++i) s[i] = lower_ascii(s[i]);
}

/* CSV column scan: restates the byte state machine of engine.c:522-605 for one partition
 * starting at byte 0 (single partition: files < 2 GiB).  file[size] must be readable (the
 * reference reads file[file_pos + 1] past quotes; pass a buffer with >= 2 zero bytes after
 * `size`).  Emits, per byte of the selected column: lower-cased text byte, file offset
 * (truncated to u32 as the reference's buffer_u32 does), quoted flag; and one '\n' sentinel
 * per row whose mapping is the offset of the delimiter that ended the field.
 * tolower() in the C locale == ASCII lower-casing.
 * Outputs must have capacity >= size + 1.  Returns the number of text bytes produced. */
uint64_t sao_csv_scan(const uint8_t* file, uint64_t size, uint32_t column_idx, uint32_t num_columns,
                      uint8_t* text, uint32_t* map, uint8_t* quoted)
{
    uint64_t pos = 0, out = 0;
    const uint64_t end_point = size;
    uint32_t col = 0;
    while (pos < end_point) {
        while (col != column_idx) {                                   /* :523-551 */
            if (file[pos] == '"') {
                ++pos;
                for (;;) {
                    if (file[pos] == '"') {
                        if (file[pos + 1] == '"') { pos += 2; if (pos >= end_point) return out; continue; }
                        ++pos; if (pos >= end_point) return out;
                        break;
                    }
                    ++pos; if (pos >= end_point) return out;
                }
            }
            if (file[pos] == ',' || file[pos] == '\n') col = (col + 1) % num_columns;
            ++pos; if (pos >= end_point) return out;
        }
        const uint8_t is_q = (file[pos] == '"');                      /* :553 */
        if (is_q) {
            ++pos;
            for (;;) {                                                /* :557-586 */
                if (file[pos] == '"' && file[pos + 1] == '"') {
                    text[out] = lower_ascii(file[pos]); map[out] = (uint32_t)pos; quoted[out] = 1; ++out;
                    pos += 2;
                    continue;
                }
                if (file[pos] == '"' && (file[pos + 1] == '\n' || file[pos + 1] == ',' || file[pos + 1] == 0)) {
                    ++pos;
                    break;
                }
                text[out] = lower_ascii(file[pos]); map[out] = (uint32_t)pos; quoted[out] = 1; ++out;
                ++pos;
                if (pos >= end_point) break;   /* unterminated quote: the reference runs off the mapping */
            }
        } else {
            for (;;) {                                                /* :588-598 */
                if (file[pos] == '\n' || file[pos] == 0 || file[pos] == ',') break;
                text[out] = lower_ascii(file[pos]); map[out] = (uint32_t)pos; quoted[out] = 0; ++out;
                ++pos;
            }
        }
        text[out] = '\n'; map[out] = (uint32_t)pos; quoted[out] = is_q; ++out;   /* :600-604 */
        ++pos;
        col = (col + 1) % num_columns;
    }
    return out;
}

/* Post-build remap, engine.c:632-651: quoted bit of rank i = quoted flag of text byte SA[i];
 * SA[i] = map[SA[i]] (SA then holds file offsets). */
void sao_csv_remap(uint32_t* sa, uint64_t n, const uint32_t* map, const uint8_t* quoted_by_text,
                   uint8_t* quoted_by_rank)
{
    /* copy_buffer_bit (engine.c:256-261) copies buffer_bit_idx / 8 whole bytes into a zeroed
     * buffer, so the flags of the last n % 8 text bytes read back as 0 in the permutation loop
     * (engine.c:632-645). */
    const uint64_t copied = n & ~(uint64_t)7;
    for (uint64_t i = 0; i < n; ++i) {
        uint32_t p = sa[i];
        quoted_by_rank[i] = (p < copied) ? quoted_by_text[p] : 0;
        sa[i] = map[p];
    }
}


/* Restates get_substring_positions_file, engine.c:920-999 -- the bisection of the reference's CSV
 * query (suffix_array.pyx:221-247 -> get_matching_records_file -> here).  `file` are the raw bytes of
 * the CSV file, sa_file the remapped array (file offsets, engine.c:647-651; one partition starting at
 * byte 0, so global_byte_start_idx == 0).  Every probe is fseek + fread of cmp_length bytes into ONE
 * buffer + tolower (:939-947): a read that crosses the end of the file fills fewer bytes and the tail
 * keeps what earlier probes left there (zero at first: the checker's reference is built with
 * zero-filled malloc, ref_prelude.h).  start / end only move on cmp == 0 (:953-957, :987-991); the
 * second loop starts at first = start (:969).  Signed bounds fence the mid == 0 wrap (as in
 * query_padded above). */
static int strncmp_u8(const uint8_t* a, const uint8_t* b, uint32_t c)
{
    for (uint32_t i = 0; i < c; ++i) {
        if (a[i] != b[i]) return (int)a[i] - (int)b[i];
        if (a[i] == 0) return 0;
    }
    return 0;
}

static void probe_file(const uint8_t* file, uint64_t fsize, uint64_t pos, uint8_t* line, uint32_t c)
{
    uint64_t avail = pos < fsize ? fsize - pos : 0;
    if (avail > c) avail = c;
    memcpy(line, file + pos, (size_t)avail);                     /* fread, :943-944 */
    for (uint32_t i = 0; i < c; ++i) line[i] = lower_ascii(line[i]);   /* :945-947 */
}

static void query_file_one(const uint8_t* file, uint64_t fsize, const uint32_t* sa, uint64_t n, uint32_t max_suffix_length,
                           const uint8_t* q, uint32_t qlen, uint32_t* out_start, uint32_t* out_end)
{
    uint32_t m = 0;
    while (m < qlen && q[m] != 0) ++m;
    const uint32_t c = m < max_suffix_length ? m : max_suffix_length;     /* :932 */
    uint32_t start = UINT32_MAX, end = UINT32_MAX;
    uint8_t* line = (uint8_t*)calloc((size_t)c + 8, 1);                   /* :934 */
    int64_t first = 0, last = (int64_t)n - 1;
    while (first <= last) {                                               /* :937-960 */
        int64_t mid = (first + last) / 2;
        probe_file(file, fsize, sa[mid], line, c);
        int cmp = strncmp_u8(line, q, c);
        if (cmp < 0) first = mid + 1;
        else { last = mid - 1; if (cmp == 0) start = (uint32_t)mid; }
    }
    if (start != UINT32_MAX) {
        first = start; last = (int64_t)n - 1;                             /* :969 */
        while (first <= last) {                                           /* :971-993 */
            int64_t mid = (first + last) / 2;
            probe_file(file, fsize, sa[mid], line, c);
            int cmp = strncmp_u8(line, q, c);
            if (cmp > 0) last = mid - 1;
            else { first = mid + 1; if (cmp == 0) end = (uint32_t)mid; }
        }
    }
    free(line);
    *out_start = start; *out_end = end;
}

int sao_query_file_batch(const uint8_t* file, uint64_t fsize, const uint32_t* sa_file, uint64_t n,
                         uint32_t max_suffix_length, const uint8_t* qbytes, const uint64_t* qoff, uint64_t nq,
                         uint32_t* start, uint32_t* end, int nthreads)
{
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(static) num_threads(nthreads)
#endif
    for (long i = 0; i < (long)nq; ++i) {
        if (n == 0) { start[i] = end[i] = UINT32_MAX; continue; }
        query_file_one(file, fsize, sa_file, n, max_suffix_length, qbytes + qoff[i], (uint32_t)(qoff[i + 1] - qoff[i]),
                       &start[i], &end[i]);
    }
    (void)nthreads;
    return 0;
}

/* Row table of a CSV file: a record ends at a '\n' with an even number of '"' in front of it
 * (RFC 4180; the line get_matching_records_file seeks for around a hit, engine.c:1217-1323, on files
 * whose quotes balance).  out[r] = first byte of row r (row 0 = header), out[rows] = size.
 * out must hold size + 2 entries.  Returns rows = record ends + 1. */
uint64_t sao_csv_row_starts(const uint8_t* file, uint64_t size, uint64_t* out)
{
    uint64_t rows = 0, quotes = 0;
    out[rows++] = 0;
    for (uint64_t i = 0; i < size; ++i) {
        if (file[i] == '"') ++quotes;
        else if (file[i] == '\n' && (quotes & 1) == 0) out[rows++] = i + 1;
    }
    out[rows] = size;
    return rows;
}

/* Segmented dedupe of hit lists (no counterpart in the reference, which keeps duplicates,
 * engine.c:1187): per query, the first hit in rank order of every record id survives. */
typedef struct { uint32_t doc, j; } sao_dj;
static int cmp_dj(const void* a, const void* b)
{
    const sao_dj* x = (const sao_dj*)a; const sao_dj* y = (const sao_dj*)b;
    if (x->doc != y->doc) return x->doc < y->doc ? -1 : 1;
    return x->j < y->j ? -1 : (x->j > y->j);
}

int sao_dedupe(const uint64_t* hit_off, uint64_t nq, const uint32_t* hit_pos, const uint32_t* hit_doc,
               uint64_t* out_off, uint32_t* out_pos, uint32_t* out_doc)
{
    uint64_t w = 0;
    for (uint64_t q = 0; q < nq; ++q) {
        const uint64_t o = hit_off[q];
        const uint32_t take = (uint32_t)(hit_off[q + 1] - o);
        out_off[q] = w;
        if (!take) continue;
        sao_dj* t = (sao_dj*)malloc((size_t)take * sizeof(sao_dj));
        uint8_t* keep = (uint8_t*)calloc(take, 1);
        if (!t || !keep) { free(t); free(keep); return -1; }
        for (uint32_t j = 0; j < take; ++j) { t[j].doc = hit_doc[o + j]; t[j].j = j; }
        qsort(t, take, sizeof(sao_dj), cmp_dj);
        for (uint32_t i = 0; i < take; ++i)
            if (i == 0 || t[i].doc != t[i - 1].doc) keep[t[i].j] = 1;
        for (uint32_t j = 0; j < take; ++j)
            if (keep[j]) { out_pos[w] = hit_pos[o + j]; out_doc[w] = hit_doc[o + j]; ++w; }
        free(t); free(keep);
    }
    out_off[nq] = w;
    return 0;
}
