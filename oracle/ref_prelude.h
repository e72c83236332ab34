/* ref_prelude.h -- force-included (-include) in front of the UNMODIFIED reference engine.c when
 * oracle/_ref/libsa_ref.so is built.  TEST INFRASTRUCTURE ONLY.
 * engine.c's parse_csv_header (engine.c:40-44, :80-85) strncpy()s column names into malloc(len)
 * blocks without a terminator and strcmp()s them later (engine.c:1425): it relies on a zero-filled
 * heap.  In a fresh process that holds; inside a long-lived Python process it does not, and the
 * column match fails.  So the reference's malloc calls get zero-filled memory. */
#include <stdlib.h>
#include <string.h>
#define malloc(n) calloc(1, (size_t)(n) + 8)
