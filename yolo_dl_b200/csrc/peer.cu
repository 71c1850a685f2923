// peer.cu — the data-parallel loss exchange as peer-memory stores over NVLink / NVSwitch, fused
// into the step (no collective kernel, no host launch per step).
//
// The reference combines the per-device loss scalars as sum_g loss_g * B_g / sum_g B_g on the
// master device every step (train/src/train/multi_gpu.rs:188-195, yolo-dl/src/loss/loss_.rs:527-577).
// With NCCL that is one 5-float all-reduce per step: a host launch and a one-CTA kernel that
// polls its peers, both per step and per rank (measured: 2-GPU step 220 us vs 205 us alone).
// Here every rank owns a MAILBOX [slots][world][8 floats] that all peers of the node map through
// CUDA IPC; the last kernel of a step writes this rank's
//     [B_g * total, B_g * iou, B_g * cls, B_g * obj, B_g, -, -, seq]
// into entry [seq % slots][rank] of EVERY peer's mailbox with plain stores (payload, then
// __threadfence_system, then the sequence number).  Capturable in a CUDA graph: the sequence
// number lives in device memory and the kernel advances it itself.  Whoever wants the combined
// loss (the logger, a few times per epoch) reads its OWN mailbox and sums the newest slot in
// which all ranks carry the same sequence number — yh_loss_collect.
#include <cstring>

#include "common.cuh"

namespace yh {

constexpr int kMailFloats = 8;
constexpr int kMaxPeers = 16;

struct PeerTab {
  float* mail[kMaxPeers];
};

__global__ void loss_publish_kernel(const float* __restrict__ losses4, const float weight,
                                    const PeerTab peers, const int world, const int rank,
                                    const int slots, unsigned* __restrict__ counter) {
  const unsigned seq = *counter + 1u;
  const int t = threadIdx.x;
  if (t < world) {
    float* dst = peers.mail[t] + ((size_t)(seq % (unsigned)slots) * world + rank) * kMailFloats;
    reinterpret_cast<volatile unsigned*>(dst)[7] = 0u;  // invalid while it is being rewritten
    __threadfence_system();
#pragma unroll
    for (int k = 0; k < 4; ++k) dst[k] = losses4[k] * weight;
    dst[4] = weight;
    __threadfence_system();  // the payload is visible to the peer before its sequence number
    reinterpret_cast<volatile unsigned*>(dst)[7] = seq;
  }
  __syncthreads();
  if (t == 0) *counter = seq;
}

}  // namespace yh

struct YhMailbox {
  int device, world, slots;
  float* mail;        // [slots][world][8] + counter
  unsigned* counter;  // behind the entries
  float* host;        // pinned copy for yh_loss_collect
  size_t floats;
};

extern "C" {

int yh_peer_mailbox_create(int device, int world, int slots, YhMailbox** out, void** mail_dev,
                           unsigned char* handle64) {
  using namespace yh;
  YH_REQUIRE(out && mail_dev && handle64, "NULL argument");
  YH_REQUIRE(world >= 1 && world <= kMaxPeers, "world=%d outside 1..%d", world, kMaxPeers);
  YH_REQUIRE(slots >= 2 && slots <= 64, "slots=%d outside 2..64", slots);
  YH_CUDA(cudaSetDevice(device));
  YhMailbox* m = new YhMailbox();
  memset(m, 0, sizeof(*m));
  m->device = device;
  m->world = world;
  m->slots = slots;
  m->floats = (size_t)slots * world * kMailFloats;
  const size_t bytes = (m->floats + 8) * sizeof(float);
  // a dedicated cudaMalloc allocation: CUDA IPC exports whole allocations
  cudaError_t e = cudaMalloc((void**)&m->mail, bytes);
  if (e == cudaSuccess) e = cudaMemset(m->mail, 0, bytes);
  if (e == cudaSuccess) e = cudaMallocHost((void**)&m->host, 2 * bytes);
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, m->mail);
  if (e != cudaSuccess) {
    set_error("mailbox allocation / IPC export failed: %s", cudaGetErrorString(e));
    cudaFree(m->mail);
    if (m->host) cudaFreeHost(m->host);
    delete m;
    cudaGetLastError();
    return YH_ERR_CUDA;
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
  memcpy(handle64, &h, 64);
  m->counter = reinterpret_cast<unsigned*>(m->mail + m->floats);
  *out = m;
  *mail_dev = m->mail;
  return YH_OK;
}

int yh_peer_mailbox_open(int device, const unsigned char* handle64, void** peer_mail) {
  using namespace yh;
  YH_REQUIRE(handle64 && peer_mail, "NULL argument");
  YH_CUDA(cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  YH_CUDA(cudaIpcOpenMemHandle(peer_mail, h, cudaIpcMemLazyEnablePeerAccess));
  return YH_OK;
}

int yh_peer_mailbox_close(int device, void* peer_mail) {
  using namespace yh;
  if (!peer_mail) return YH_OK;
  YH_CUDA(cudaSetDevice(device));
  YH_CUDA(cudaIpcCloseMemHandle(peer_mail));
  return YH_OK;
}

int yh_peer_mailbox_destroy(YhMailbox* m) {
  if (!m) return YH_OK;
  cudaSetDevice(m->device);
  cudaFree(m->mail);
  if (m->host) cudaFreeHost(m->host);
  delete m;
  return YH_OK;
}

int yh_loss_publish(YhMailbox* m, const float* losses4_dev, float weight, void* const* peer_mails,
                    int rank, void* stream) {
  using namespace yh;
  YH_REQUIRE(m && losses4_dev && peer_mails, "NULL argument");
  YH_REQUIRE(rank >= 0 && rank < m->world, "rank=%d outside 0..%d", rank, m->world - 1);
  PeerTab tab;
  memset(&tab, 0, sizeof(tab));
  for (int r = 0; r < m->world; ++r) {
    YH_REQUIRE(peer_mails[r] != nullptr, "peer_mails[%d] is NULL", r);
    tab.mail[r] = (float*)peer_mails[r];
  }
  YH_REQUIRE(peer_mails[rank] == (void*)m->mail, "peer_mails[rank] must be this rank's own mailbox");
  cudaStream_t s = (cudaStream_t)stream;
  YH_CUDA(cudaSetDevice(m->device));
  YH_TIME(YH_K_LOSS_PUBLISH, s);
  loss_publish_kernel<<<1, 32, 0, s>>>(losses4_dev, weight, tab, m->world, rank, m->slots,
                                       m->counter);
  YH_LAUNCHED(s);
  return YH_OK;
}

int yh_loss_collect(YhMailbox* m, float* out4, unsigned* out_seq, void* stream) {
  using namespace yh;
  YH_REQUIRE(m && out4 && out_seq, "NULL argument");
  cudaStream_t s = (cudaStream_t)stream;
  YH_CUDA(cudaSetDevice(m->device));
  // two copies: a slot is used only if both show the same bytes (peers may be writing; an entry
  // is invalidated before and stamped after its payload)
  float* second = m->host + m->floats + 8;
  YH_CUDA(cudaMemcpyAsync(m->host, m->mail, m->floats * sizeof(float), cudaMemcpyDeviceToHost, s));
  YH_CUDA(cudaMemcpyAsync(second, m->mail, m->floats * sizeof(float), cudaMemcpyDeviceToHost, s));
  YH_CUDA(cudaStreamSynchronize(s));
  // newest slot in which every rank carries the same sequence number
  unsigned best = 0;
  int best_slot = -1;
  for (int slot = 0; slot < m->slots; ++slot) {
    const float* e0 = m->host + (size_t)slot * m->world * kMailFloats;
    unsigned seq;
    memcpy(&seq, e0 + 7, 4);
    bool same = seq != 0;
    for (int r = 1; r < m->world && same; ++r) {
      unsigned q;
      memcpy(&q, e0 + (size_t)r * kMailFloats + 7, 4);
      same = q == seq;
    }
    same = same && memcmp(e0, second + (size_t)slot * m->world * kMailFloats,
                          (size_t)m->world * kMailFloats * sizeof(float)) == 0;
    if (same && seq > best) {
      best = seq;
      best_slot = slot;
    }
  }
  *out_seq = best;
  out4[0] = out4[1] = out4[2] = out4[3] = 0.f;
  if (best_slot < 0) return YH_OK;  // nothing complete yet: seq 0
  // sum in rank order, f32 like the reference's tensor arithmetic (tch-goodies/src/tensor.rs:66-80)
  float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
  for (int r = 0; r < m->world; ++r) {
    const float* e = m->host + ((size_t)best_slot * m->world + r) * kMailFloats;
    for (int k = 0; k < 5; ++k) acc[k] += e[k];
  }
  for (int k = 0; k < 4; ++k) out4[k] = acc[k] / acc[4];
  return YH_OK;
}

}  // extern "C"
