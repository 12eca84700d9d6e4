// Argument blocks of the sparse (map container) collectives on the GPU.
//
// A GPU map is a pair of device arrays (int64 keys, T values). String keys are hashed to 64 bits
// on the host (binding side); the reference's maps are Map<String,T> (PCS:2053-2081).
// Ownership of a key (the reference: String.hashCode() % slaveNum, PCS:2063-2071):
//   MP4J_MAP_PART_MIX    owner = splitmix64(key) % p                       (integer ids)
//   MP4J_MAP_PART_JAVA32 owner = ((int32)key) mod p, negatives wrapped     (low 32 bits carry
//                        java.lang.String.hashCode, so placement matches the reference)
//   MP4J_MAP_PART_TAGGED owner = top 4 bits of the key (the caller addresses a rank explicitly)
#pragma once
#include "peer.cuh"

namespace mp4j {

constexpr long long kMapEmpty = (long long)0x8000000000000000ull;  // reserved key: empty slot

enum MapPartition { MP4J_MAP_PART_MIX = 0, MP4J_MAP_PART_JAVA32 = 1, MP4J_MAP_PART_TAGGED = 2 };

struct MapExchangeArgs {
  PeerTable tab;
  int partition;
  int pad_;
  long long cap;                      // pairs per (destination, source) receive region
  unsigned long long rkeys_off;       // heap: int64 [p regions][cap]   (region index = source rank)
  unsigned long long rvals_off;       // heap: T     [p regions][cap]
  unsigned long long rcnt_off;        // heap: uint64 [p]               (pairs received from each source)
  unsigned long long cursor_off;      // heap (local use): uint64 [p] write cursors, uint64 overflow flag at [p]
  const long long* keys[kMaxRanks];   // local inputs, indexed by rank in emulation, [0] otherwise
  const void* vals[kMaxRanks];
  long long n[kMaxRanks];
};

struct MapTable {
  long long* keys;      // [size] open addressing, kMapEmpty = free
  void* vals;           // [size]
  unsigned long long mask;  // size - 1 (size is a power of two)
};

}  // namespace mp4j
