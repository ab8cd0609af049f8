/* xxh64_ref.c — C restatement of canonical XXH64 (TEST INFRASTRUCTURE; see oracle/xxh64.py for provenance:
 * third-party xxhash-rust 0.8, call site reference src/engine/block_manager.rs:109-123).
 * Built by oracle/Makefile into oracle/_build/liboracle.so and cross-checked against the Python restatement,
 * the `xxhash` package and tests/golden/xxh64.json. */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#define P1 0x9E3779B185EBCA87ULL
#define P2 0xC2B2AE3D27D4EB4FULL
#define P3 0x165667B19E3779F9ULL
#define P4 0x85EBCA77C2B2AE63ULL
#define P5 0x27D4EB2F165667C5ULL

static uint64_t rotl(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
static uint64_t rd64(const uint8_t* p) { uint64_t v; memcpy(&v, p, 8); return v; } /* little-endian host */
static uint32_t rd32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
static uint64_t round64(uint64_t acc, uint64_t in) { acc += in * P2; return rotl(acc, 31) * P1; }
static uint64_t merge(uint64_t acc, uint64_t v) { acc ^= round64(0, v); return acc * P1 + P4; }

uint64_t oracle_xxh64(const uint8_t* p, size_t n, uint64_t seed) {
  const uint8_t* end = p + n;
  uint64_t h;
  if (n >= 32) {
    uint64_t v1 = seed + P1 + P2, v2 = seed + P2, v3 = seed, v4 = seed - P1;
    do {
      v1 = round64(v1, rd64(p)); v2 = round64(v2, rd64(p + 8));
      v3 = round64(v3, rd64(p + 16)); v4 = round64(v4, rd64(p + 24));
      p += 32;
    } while (p + 32 <= end);
    h = rotl(v1, 1) + rotl(v2, 7) + rotl(v3, 12) + rotl(v4, 18);
    h = merge(h, v1); h = merge(h, v2); h = merge(h, v3); h = merge(h, v4);
  } else {
    h = seed + P5;
  }
  h += (uint64_t)n;
  while (p + 8 <= end) { h ^= round64(0, rd64(p)); h = rotl(h, 27) * P1 + P4; p += 8; }
  if (p + 4 <= end) { h ^= (uint64_t)rd32(p) * P1; h = rotl(h, 23) * P2 + P3; p += 4; }
  while (p < end) { h ^= (*p) * P5; h = rotl(h, 11) * P1; ++p; }
  h ^= h >> 33; h *= P2; h ^= h >> 29; h *= P3; h ^= h >> 32;
  return h;
}

/* BlockManager::compute_hash (block_manager.rs:109-123): [prefix u64 LE]? || tokens i64 LE, seed 0. */
uint64_t oracle_compute_hash(const int64_t* tokens, size_t n, int has_prefix, uint64_t prefix) {
  uint8_t stack[8 + 8 * 512];
  uint8_t* buf = stack;
  uint8_t* heap = NULL;
  size_t bytes = (has_prefix ? 8 : 0) + 8 * n;
  if (bytes > sizeof(stack)) { extern void* malloc(size_t); heap = (uint8_t*)malloc(bytes); buf = heap; }
  size_t o = 0;
  if (has_prefix) { memcpy(buf, &prefix, 8); o = 8; }
  memcpy(buf + o, tokens, 8 * n);
  uint64_t h = oracle_xxh64(buf, bytes, 0);
  if (heap) { extern void free(void*); free(heap); }
  return h;
}
