// ce_twoshot.cuh -- the copy-engine two-shot allreduce as a reusable host-side engine.
//
// Measured on this fabric (profiles/r02_fabric.md): loads and stores issued by SMs cross NVLink as 128-byte packets
// (1.125 / 1.1875 link bytes per payload byte; 667-700 GB/s of payload per direction with both directions busy), while
// the GPUs' DMA engines move 770 GB/s per direction in the same situation. This engine lets the copy engines do every
// transfer and keeps only the arithmetic in a kernel:
//   scatter  rank r copies its words of slice q into a staging slot in q's memory      (n-1 peer copies per chunk)
//   add      rank q adds the n-1 staged copies to its own words of slice q             (ce_add_staged_kernel, local HBM)
//   gather   rank q copies the reduced slice q into every rank's buffer                (n-1 peer copies per chunk)
// Slices are cut into chunks so that the add of chunk k runs while the engines move chunk k+1 (scatter) and k-1
// (gather). Ordering is by CUDA events across the GPUs' streams (one process drives all GPUs); no kernel ever waits on
// another GPU. Peers are visited in rotated order so that every GPU receives from one sender at a time.
//
// Used by gpuprobe_allreduce_i32_ce (allreduce.cu, the whole buffer) and by the NVLS + copy-engine hybrid (nvls.cu, the
// tail of the buffer while the multimem kernel reduces the head).
#pragma once
#include "common.cuh"

namespace gp {

// x[i] += sum over the staged copies of the same words. All operands are in this GPU's HBM; 128-bit accesses.
static __global__ void __launch_bounds__(256)
ce_add_staged_kernel(uint32_t *__restrict__ x, const uint32_t *__restrict__ stage, uint64_t slot_stride, int nslots, uint64_t n) {
    const uint64_t tid = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    const uint64_t n4 = n / 4;  // x, stage and slot_stride are 16-byte aligned (chunks start on 32-word boundaries)
    for (uint64_t v = tid; v < n4; v += stride) {
        uint4 acc = *reinterpret_cast<const uint4 *>(x + v * 4);
        for (int s = 0; s < nslots; ++s) {
            const uint4 t = ldg128_stream(stage + static_cast<uint64_t>(s) * slot_stride + v * 4);
            acc.x += t.x;
            acc.y += t.y;
            acc.z += t.z;
            acc.w += t.w;
        }
        *reinterpret_cast<uint4 *>(x + v * 4) = acc;
    }
    for (uint64_t i = n4 * 4 + tid; i < n; i += stride) {
        uint32_t acc = x[i];
        for (int s = 0; s < nslots; ++s) acc += stage[static_cast<uint64_t>(s) * slot_stride + i];
        x[i] = acc;
    }
}

struct CeRank {
    int ord = -1;              // CUDA ordinal
    int sm_count = 0;
    uint32_t *buf = nullptr;   // this rank's `count` words (16-byte aligned), reduced in place
    uint32_t *stage = nullptr; // CeTwoShot::stage_words(count, n) words of staging in this rank's memory
    cudaStream_t main = nullptr;  // the round is ordered after what `main` holds at issue time, and `main` ends up waiting for it
};

class CeTwoShot {
public:
    static uint64_t slice_words(uint64_t count, int nd) {
        const uint64_t per = (count + nd - 1) / nd;
        return (per + 31) & ~31ull;
    }
    static uint64_t stage_words(uint64_t count, int nd) { return static_cast<uint64_t>(nd - 1) * slice_words(count, nd); }

    // The devices must already have peer access to one another. `chunks` <= 0 picks the measured default.
    cudaError_t init(const std::vector<CeRank> &ranks, uint64_t count, int chunks) {
        R_ = ranks;
        nd_ = static_cast<int>(ranks.size());
        count_ = count;
        per_ = slice_words(count, nd_);
        // measured at 2 ranks, 1 GiB: 1 chunk 635, 2 chunks 748, 4 731, 8 699, 16 647 GB/s busbw
        chunks_ = chunks > 0 ? chunks : 2;
        if (const char *e = getenv("GPUPROBE_CE_CHUNKS")) chunks_ = std::max(1, std::min(64, atoi(e)));
        while (chunks_ > 1 && per_ / chunks_ < (1u << 18)) chunks_ /= 2;
        chunk_words_ = ((per_ + chunks_ - 1) / chunks_ + 31) & ~31ull;
        S_.resize(nd_);
        cudaError_t e;
        for (int r = 0; r < nd_; ++r) {
            if ((e = cudaSetDevice(R_[r].ord)) != cudaSuccess) return e;
            Side &s = S_[r];
            int lo = 0, hi = 0;
            if ((e = cudaDeviceGetStreamPriorityRange(&lo, &hi)) != cudaSuccess) return e;
            if ((e = cudaStreamCreateWithFlags(&s.scatter, cudaStreamNonBlocking)) != cudaSuccess) return e;
            if ((e = cudaStreamCreateWithFlags(&s.gather, cudaStreamNonBlocking)) != cudaSuccess) return e;
            // the add must not queue behind a long-running kernel of the caller on the same GPU
            if ((e = cudaStreamCreateWithPriority(&s.add, cudaStreamNonBlocking, hi)) != cudaSuccess) return e;
            if ((e = cudaEventCreateWithFlags(&s.go, cudaEventDisableTiming)) != cudaSuccess) return e;
            if ((e = cudaEventCreateWithFlags(&s.gathered, cudaEventDisableTiming)) != cudaSuccess) return e;
            s.scattered.assign(chunks_, nullptr);
            s.added.assign(chunks_, nullptr);
            for (auto &ev : s.scattered)
                if ((e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)) != cudaSuccess) return e;
            for (auto &ev : s.added)
                if ((e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)) != cudaSuccess) return e;
        }
        return cudaSuccess;
    }

    // One in-place round, issued asynchronously. Afterwards every rank's `main` stream waits for the whole round
    // (every gather, incoming and outgoing), so consecutive rounds and whatever the caller enqueues next are ordered.
    cudaError_t issue_round() {
        cudaError_t e;
        for (int r = 0; r < nd_; ++r) {
            if ((e = cudaSetDevice(R_[r].ord)) != cudaSuccess) return e;
            if ((e = cudaEventRecord(S_[r].go, R_[r].main)) != cudaSuccess) return e;
            if ((e = cudaStreamWaitEvent(S_[r].scatter, S_[r].go, 0)) != cudaSuccess) return e;
            if ((e = cudaStreamWaitEvent(S_[r].gather, S_[r].go, 0)) != cudaSuccess) return e;
            if ((e = cudaStreamWaitEvent(S_[r].add, S_[r].go, 0)) != cudaSuccess) return e;
        }
        for (int k = 0; k < chunks_; ++k)  // scatter: chunk by chunk, every rank, peers in rotated order
            for (int r = 0; r < nd_; ++r) {
                if ((e = cudaSetDevice(R_[r].ord)) != cudaSuccess) return e;
                for (int j = 1; j < nd_; ++j) {
                    const int q = (r + j) % nd_;
                    uint64_t off, len;
                    chunk_range(slice_len(q), k, &off, &len);
                    if (!len) continue;
                    const int slot = r < q ? r : r - 1;
                    e = cudaMemcpyPeerAsync(R_[q].stage + static_cast<uint64_t>(slot) * per_ + off, R_[q].ord, R_[r].buf + slice_lo(q) + off,
                                            R_[r].ord, len * sizeof(uint32_t), S_[r].scatter);
                    if (e != cudaSuccess) return e;
                }
                if ((e = cudaEventRecord(S_[r].scattered[k], S_[r].scatter)) != cudaSuccess) return e;
            }
        for (int k = 0; k < chunks_; ++k)  // add, then gather the reduced chunk to everyone
            for (int q = 0; q < nd_; ++q) {
                if ((e = cudaSetDevice(R_[q].ord)) != cudaSuccess) return e;
                for (int r = 0; r < nd_; ++r)
                    if (r != q && (e = cudaStreamWaitEvent(S_[q].add, S_[r].scattered[k], 0)) != cudaSuccess) return e;
                uint64_t off, len;
                chunk_range(slice_len(q), k, &off, &len);
                if (len) {
                    const uint64_t want = (len / 4 + 255) / 256;
                    const int grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(want, static_cast<uint64_t>(R_[q].sm_count) * 8)));
                    ce_add_staged_kernel<<<grid, 256, 0, S_[q].add>>>(R_[q].buf + slice_lo(q) + off, R_[q].stage + off, per_, nd_ - 1, len);
                    if ((e = cudaGetLastError()) != cudaSuccess) return e;
                }
                if ((e = cudaEventRecord(S_[q].added[k], S_[q].add)) != cudaSuccess) return e;
                if ((e = cudaStreamWaitEvent(S_[q].gather, S_[q].added[k], 0)) != cudaSuccess) return e;
                for (int j = 1; j < nd_ && len; ++j) {
                    const int p = (q + j) % nd_;
                    e = cudaMemcpyPeerAsync(R_[p].buf + slice_lo(q) + off, R_[p].ord, R_[q].buf + slice_lo(q) + off, R_[q].ord,
                                            len * sizeof(uint32_t), S_[q].gather);
                    if (e != cudaSuccess) return e;
                }
            }
        for (int r = 0; r < nd_; ++r) {
            if ((e = cudaSetDevice(R_[r].ord)) != cudaSuccess) return e;
            if ((e = cudaEventRecord(S_[r].gathered, S_[r].gather)) != cudaSuccess) return e;
        }
        for (int r = 0; r < nd_; ++r) {
            if ((e = cudaSetDevice(R_[r].ord)) != cudaSuccess) return e;
            for (int x = 0; x < nd_; ++x)
                if ((e = cudaStreamWaitEvent(R_[r].main, S_[x].gathered, 0)) != cudaSuccess) return e;
        }
        return cudaSuccess;
    }

    // The caller has synchronised the devices.
    void destroy() {
        for (int r = 0; r < static_cast<int>(S_.size()); ++r) {
            if (cudaSetDevice(R_[r].ord) != cudaSuccess) continue;
            Side &s = S_[r];
            if (s.scatter) cudaStreamDestroy(s.scatter);
            if (s.gather) cudaStreamDestroy(s.gather);
            if (s.add) cudaStreamDestroy(s.add);
            if (s.go) cudaEventDestroy(s.go);
            if (s.gathered) cudaEventDestroy(s.gathered);
            for (cudaEvent_t ev : s.scattered)
                if (ev) cudaEventDestroy(ev);
            for (cudaEvent_t ev : s.added)
                if (ev) cudaEventDestroy(ev);
        }
        S_.clear();
        (void)cudaGetLastError();
    }

    int chunks() const { return chunks_; }

private:
    struct Side {
        cudaStream_t scatter = nullptr, gather = nullptr, add = nullptr;
        cudaEvent_t go = nullptr, gathered = nullptr;
        std::vector<cudaEvent_t> scattered, added;  // per chunk
    };
    uint64_t slice_lo(int q) const { return std::min<uint64_t>(per_ * static_cast<uint64_t>(q), count_); }
    uint64_t slice_len(int q) const { return std::min<uint64_t>(slice_lo(q) + per_, count_) - slice_lo(q); }
    // words [off, off+len) of chunk k inside a slice of `len_slice` words
    void chunk_range(uint64_t len_slice, int k, uint64_t *off, uint64_t *len) const {
        *off = std::min<uint64_t>(chunk_words_ * static_cast<uint64_t>(k), len_slice);
        *len = std::min<uint64_t>(*off + chunk_words_, len_slice) - *off;
    }
    std::vector<CeRank> R_;
    std::vector<Side> S_;
    int nd_ = 0, chunks_ = 2;
    uint64_t count_ = 0, per_ = 0, chunk_words_ = 0;
};

}  // namespace gp
