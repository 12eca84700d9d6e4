// Collectives on OPAQUE payloads (String / Object operands, maps whose values the core cannot reduce): the
// data plane - ring reduce-scatter, ring allgather, binomial gather and scatter (PCS:1329-1373, 694-737,
// 440-520, 975-1050) - runs here; the binding supplies only the codec through two callbacks:
//
//   get(user, part, &data, &n)   serialized bytes of local part `part` (must stay valid until the next callback)
//   put(user, part, data, n)     deserialize; merge into local part `part` (reduce-scatter) or store it (others)
//
// That is the reference's operand model (ObjectOperand.java:114-156: the operand serializes, the collective
// moves bytes, recv-and-merge applies the operator per element) with the collective in C++ and the operand in
// the binding's language. Inner nodes of the gather / scatter trees forward bundles of peers' payloads as raw
// bytes - they are never deserialized on the way.
#include <string.h>

#include <map>
#include <string>
#include <vector>

#include "process_comm.h"

namespace mp4j {

namespace {

// bundle = [u32 count] { [i32 rank] [u64 len] [bytes] } *
void bundle_add(std::string* b, int rank, const char* data, size_t n) {
  if (b->empty()) b->assign(4, '\0');
  uint32_t cnt;
  memcpy(&cnt, b->data(), 4);
  ++cnt;
  memcpy(&(*b)[0], &cnt, 4);
  const int32_t r = rank;
  const uint64_t len = n;
  b->append((const char*)&r, 4);
  b->append((const char*)&len, 8);
  b->append(data, n);
}

template <typename F>
void bundle_each(const std::string& b, F&& fn) {
  if (b.size() < 4) return;
  uint32_t cnt;
  memcpy(&cnt, b.data(), 4);
  size_t off = 4;
  for (uint32_t i = 0; i < cnt; ++i) {
    if (off + 12 > b.size()) throw Error("blob collective: truncated bundle");
    int32_t r;
    uint64_t len;
    memcpy(&r, b.data() + off, 4);
    memcpy(&len, b.data() + off + 4, 8);
    off += 12;
    if (off + len > b.size()) throw Error("blob collective: malformed bundle");
    fn((int)r, b.data() + off, (size_t)len);
    off += len;
  }
}

void call_get(ProcessComm::BlobGet get, void* user, int part, const char** d, size_t* n) {
  *d = nullptr;
  *n = 0;
  if (get(user, part, d, n) != 0) throw Error("blob collective: the binding's serializer failed");
}
void call_put(ProcessComm::BlobPut put, void* user, int part, const char* d, size_t n) {
  if (put(user, part, d, n) != 0) throw Error("blob collective: the binding's deserializer / reducer failed");
}

}  // namespace

void ProcessComm::blob_reduce_scatter(BlobGet get, BlobPut put, void* user) {
  const int p = n_;
  if (p == 1) return;
  std::lock_guard<std::mutex> lk(op_mu_);
  const int next = (rank_ + 1) % p, prev = (rank_ - 1 + p) % p;
  for (int s = 0; s < p - 1; ++s) {  // step s sends part (rank-1-s), merges what arrives into part (rank-2-s)
    const int sseg = ((rank_ - 1 - s) % p + p) % p, rseg = ((rank_ - 2 - s) % p + p) % p;
    const char* d;
    size_t n;
    call_get(get, user, sseg, &d, &n);
    const std::string in = sendrecv_bytes(next, d, n, prev);
    call_put(put, user, rseg, in.data(), in.size());
  }
}

void ProcessComm::blob_allgather(BlobGet get, BlobPut put, void* user) {
  const int p = n_;
  if (p == 1) return;
  std::lock_guard<std::mutex> lk(op_mu_);
  const int next = (rank_ + 1) % p, prev = (rank_ - 1 + p) % p;
  std::vector<std::string> blocks((size_t)p);
  {
    const char* d;
    size_t n;
    call_get(get, user, rank_, &d, &n);
    blocks[(size_t)rank_].assign(d, n);
  }
  for (int s = 0; s < p - 1; ++s) {
    const int sseg = ((rank_ - s) % p + p) % p, rseg = ((rank_ - 1 - s) % p + p) % p;
    blocks[(size_t)rseg] = sendrecv_bytes(next, blocks[(size_t)sseg].data(), blocks[(size_t)sseg].size(), prev);
    call_put(put, user, rseg, blocks[(size_t)rseg].data(), blocks[(size_t)rseg].size());
    if (sseg != rank_) std::string().swap(blocks[(size_t)sseg]);  // forwarded: no longer needed
  }
}

void ProcessComm::blob_gather(int root, BlobGet get, BlobPut put, void* user) {
  const int p = n_;
  if (root < 0 || root >= p) throw Error("root rank out of range");
  if (p == 1) return;
  std::lock_guard<std::mutex> lk(op_mu_);
  const int vr = ((rank_ - root) % p + p) % p;
  std::string acc;
  {
    const char* d;
    size_t n;
    call_get(get, user, rank_, &d, &n);
    bundle_add(&acc, rank_, d, n);
  }
  for (int m = 1; m < p; m <<= 1) {
    if (vr & m) {
      send_bytes(((vr - m) + root) % p, acc.data(), acc.size());
      return;
    }
    if (vr + m < p) {
      const std::string in = recv_bytes(((vr + m) + root) % p);
      bundle_each(in, [&](int r, const char* d, size_t n) { bundle_add(&acc, r, d, n); });
    }
  }
  // root: everything is here
  bundle_each(acc, [&](int r, const char* d, size_t n) {
    if (r != rank_) call_put(put, user, r, d, n);
  });
}

void ProcessComm::blob_scatter(int root, BlobGet get, BlobPut put, void* user) {
  const int p = n_;
  if (root < 0 || root >= p) throw Error("root rank out of range");
  if (p == 1) return;
  std::lock_guard<std::mutex> lk(op_mu_);
  const int vr = ((rank_ - root) % p + p) % p;
  int top = 1;
  while (top < p) top <<= 1;
  top >>= 1;
  std::map<int, std::string> acc;  // rank -> payload bytes this node still has to deliver (or keep)
  int m;
  if (vr == 0) {
    for (int q = 0; q < p; ++q) {
      if (q == rank_) continue;
      const char* d;
      size_t n;
      call_get(get, user, q, &d, &n);
      acc[q].assign(d, n);
    }
    m = top;
  } else {
    const int lsb = vr & -vr;
    const std::string in = recv_bytes(((vr - lsb) + root) % p);
    bundle_each(in, [&](int r, const char* d, size_t n) { acc[r].assign(d, n); });
    m = lsb >> 1;
  }
  for (; m >= 1; m >>= 1) {
    if (vr + m >= p) continue;
    std::string out;
    for (int v = vr + m; v < std::min(vr + 2 * m, p); ++v) {
      const int r = (v + root) % p;
      auto it = acc.find(r);
      if (it == acc.end()) throw Error("blob scatter: payload missing in the tree");
      bundle_add(&out, r, it->second.data(), it->second.size());
      acc.erase(it);
    }
    send_bytes(((vr + m) + root) % p, out.data(), out.size());
  }
  if (vr != 0) {
    auto it = acc.find(rank_);
    if (it == acc.end()) throw Error("blob scatter: own payload missing");
    call_put(put, user, rank_, it->second.data(), it->second.size());
  }
}

}  // namespace mp4j
