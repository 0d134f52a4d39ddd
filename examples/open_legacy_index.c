/* Minimal C host program over include/lance_b200.h: parses a legacy IVF_PQ index.idx with the library's native reader
 * and, when a CUDA device is present, uploads it and runs one search.  Plain C11 -- the boundary carries no C++ or torch
 * types.  Build:  gcc -std=c11 -Iinclude examples/open_legacy_index.c -Llance_b200 -llance_b200 -o open_legacy_index
 * Run:    LD_LIBRARY_PATH=lance_b200 ./open_legacy_index <path/to/index.idx> */
#include <stdio.h>
#include <stdlib.h>

#include "lance_b200.h"

int main(int argc, char** argv) {
  if (argc < 2) {
    fprintf(stderr, "usage: %s index.idx\n", argv[0]);
    return 2;
  }
  printf("%s\n", lb200_version());
  lb200_index_file* f = NULL;
  if (lb200_index_file_open(argv[1], &f) != LB200_OK) {
    fprintf(stderr, "open failed: %s\n", lb200_last_error());
    return 1;
  }
  lb200_index_file_info info;
  lb200_index_file_get_info(f, &info);
  printf("name=%s column=%s metric=%d dim=%u nlist=%u M=%u bits=%u rows=%llu first_row_id=%llu\n", info.name, info.column,
         info.metric, info.dim, info.nlist, info.num_sub_vectors, info.num_bits, (unsigned long long)info.num_rows,
         info.num_rows ? (unsigned long long)info.row_ids[0] : 0ull);
  lb200_index* ix = NULL;
  int rc = lb200_index_open(0, argv[1], NULL, &ix);
  if (rc == LB200_ECUDA) {
    printf("no CUDA device: parsed only (%s)\n", lb200_last_error());
  } else if (rc != LB200_OK) {
    fprintf(stderr, "index_open failed: %s\n", lb200_last_error());
    lb200_index_file_close(f);
    return 1;
  } else {
    float* q = (float*)calloc(info.dim, sizeof(float)); /* the all-zero key of rust/lance/src/dataset.rs:4326-4329 */
    float dist[10];
    uint64_t rid[10];
    uint32_t cnt = 0;
    rc = lb200_search(ix, q, 1, 4, 10, 0, NULL, NULL, dist, rid, &cnt);
    if (rc != LB200_OK) fprintf(stderr, "search failed: %s\n", lb200_last_error());
    for (uint32_t i = 0; i < cnt; i++) printf("  %u: row %llu  distance %.6f\n", i, (unsigned long long)rid[i], dist[i]);
    free(q);
    lb200_index_destroy(ix);
  }
  lb200_index_file_close(f);
  return rc == LB200_OK || rc == LB200_ECUDA ? 0 : 1;
}
