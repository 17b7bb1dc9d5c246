// Host-side runtime glue of libddpb200: error strings, driver entry points and
// device queries.  No counterpart in the reference (pure Python).
#include <stdarg.h>
#include <string.h>

#include "common.cuh"
#include <stdlib.h>

namespace ddpb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

const char* last_error() { return g_err; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || p == nullptr) {
      set_error("cuTensorMapEncodeTiled is not available from this driver (%s)", cudaGetErrorString(e));
      return nullptr;
    }
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

int encode_tmap(CUtensorMap* map, CUtensorMapDataType dtype, uint32_t rank, const void* base,
                const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                CUtensorMapSwizzle swizzle, CUtensorMapL2promotion l2promo) {
  EncodeTiledFn fn = get_encode_fn();
  if (fn == nullptr) return 1;
  cuuint64_t gdims[5];
  cuuint64_t gstrides[5];
  cuuint32_t gbox[5];
  cuuint32_t estr[5];
  for (uint32_t i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = 1;
    if (i + 1 < rank) gstrides[i] = strides_bytes[i];
  }
  CUresult r = fn(map, dtype, rank, const_cast<void*>(base), gdims, gstrides, gbox, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, l2promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %u, base %p, dims %llu/%llu, box %u/%u)",
              static_cast<int>(r), rank, base, (unsigned long long)gdims[0], (unsigned long long)gdims[1],
              gbox[0], gbox[1]);
    return 1;
  }
  return 0;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

static int g_pdl = -1;   // -1: not decided yet (environment DDPB_PDL, default on)

bool pdl_enabled() {
  if (g_pdl < 0) {
    const char* e = getenv("DDPB_PDL");
    g_pdl = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return g_pdl != 0;
}

}  // namespace ddpb

namespace ddpb {
__global__ void stamp_kernel(unsigned long long* slot) { *slot = global_timer_ns(); }
}  // namespace ddpb

// Diagnostics: writes %globaltimer (ns) to *slot when the stream reaches this point -- a one-thread
// kernel, launched without programmatic overlap so it runs after everything before it completed.
extern "C" int ddpb_stamp(unsigned long long* slot, cudaStream_t stream) {
  ddpb::stamp_kernel<<<1, 1, 0, stream>>>(slot);
  DDPB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

namespace ddpb {
// One thread: bump a device-side sequence number and store {sequence, value bits} as ONE 8-byte word
// into mapped pinned host memory.  The host polls that word: when the sequence number it expects
// shows up, the value next to it is the one written with it (an aligned 8-byte store reaches host
// memory as one transaction).  Lets a training loop read every step's loss without a D2H copy call
// or a stream synchronisation, and without waiting for the rest of the step (the loss is known
// after the classifier, long before backward ends).
__global__ void publish_scalar_kernel(const float* __restrict__ value, unsigned int* __restrict__ dev_seq,
                                      unsigned long long* host_slot) {
  const unsigned int seq = *dev_seq + 1u;
  *dev_seq = seq;
  const unsigned long long word =
      (static_cast<unsigned long long>(seq) << 32) | static_cast<unsigned long long>(__float_as_uint(*value));
  asm volatile("st.global.relaxed.sys.u64 [%0], %1;" ::"l"(host_slot), "l"(word) : "memory");
  __threadfence_system();
}
}  // namespace ddpb

// value: device fp32 scalar; dev_seq: device counter (starts at 0, owned by the caller); host_slot:
// 8-byte aligned word in pinned (mapped) host memory, device-accessible address.  After this point of
// the stream has executed, *host_slot == (new sequence number << 32) | bits of *value.
extern "C" int ddpb_publish_scalar(const float* value, unsigned int* dev_seq, unsigned long long* host_slot,
                                   cudaStream_t stream) {
  DDPB_REQUIRE(value != nullptr && dev_seq != nullptr && host_slot != nullptr, "publish_scalar: null argument");
  DDPB_REQUIRE(reinterpret_cast<uintptr_t>(host_slot) % 8 == 0, "publish_scalar: host slot must be 8-byte aligned");
  ddpb::publish_scalar_kernel<<<1, 1, 0, stream>>>(value, dev_seq, host_slot);
  DDPB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Programmatic dependent launch between consecutive kernels: 1 on (default), 0 off.
extern "C" void ddpb_set_pdl(int on) { ddpb::g_pdl = on ? 1 : 0; }
extern "C" int ddpb_get_pdl() { return ddpb::pdl_enabled() ? 1 : 0; }

extern "C" const char* ddpb_last_error() { return ddpb::last_error(); }

extern "C" int ddpb_sm_count() { return ddpb::sm_count(); }

// Library/ABI version: bump when an entry point's signature changes.
extern "C" int ddpb_abi_version() { return 5; }
