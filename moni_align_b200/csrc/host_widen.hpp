// Host side of the u64 result columns (moni_gpu_mems): records leave the device in their narrowest form (2 / 4 / 8 bytes each)
// and a small pool of host threads widens them into the caller's u64 arrays with streaming stores, chunk by chunk, while the
// GPU works on the next chunks.  A format conversion of results the GPU computed (like moni_gpu_ms_runs_expand), not a compute
// path: the reference's record type is std::pair<size_t, size_t> (src/mems.cpp:277-279), 16 bytes per MEM, and the PCIe link
// -- not the kernels -- bounded the call when those 16 bytes crossed it.
#pragma once
#include <cstddef>
#include <cstdint>
#include <functional>

namespace moni {

// cfmt: 1 = u16 records (query offset << 8) | length; 2 = u32 records (query offset << 16) | length;
//       4 = two u32 columns (src = query offsets, src2 = lengths).  Writes dpos[0..k), dlen[0..k).
void widen_records(int cfmt, const void* src, const void* src2, uint64_t k, uint64_t* dpos, uint64_t* dlen);

// moni_gpu_ms: the 7-byte-per-base form of moni_gpu_ms_packed (pointer low word, pointer bits 32..39, u16 length) to the two u64
// vectors.  A 40-bit value >= 2^40 - 2^31 is one of the reference's wrapped pointers 2^64 - j (include/moni_gpu.h) and is
// widened as such.
void widen_ms(const uint32_t* lo, const uint8_t* hi, const uint16_t* len, uint64_t k, uint64_t* dptr, uint64_t* dlen);

// Process-wide pool.  threads() == 0: widening is off (too few CPUs allowed, or option "widen_threads" = 0).
class WidenPool {
  public:
    static WidenPool& get();
    int threads() const { return n_; }
    void resize(int n);                                        // joins the old workers; n = 0 switches the pool off
    // fn(slice) for slice in [0, n_slices) on the pool's threads; returns when all are done.  One parallel_for at a time.
    void parallel_for(int n_slices, const std::function<void(int)>& fn);
    // widen_records split over the pool
    void widen(int cfmt, const void* src, const void* src2, uint64_t k, uint64_t* dpos, uint64_t* dlen);
    void widen_ms(const uint32_t* lo, const uint8_t* hi, const uint16_t* len, uint64_t k, uint64_t* dptr, uint64_t* dlen);
    static int default_threads();                              // from the CPUs this process may run on

  private:
    WidenPool();
    ~WidenPool();
    struct Impl;
    Impl* impl_;
    int n_ = 0;
};

}  // namespace moni
