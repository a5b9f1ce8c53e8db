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
// another GPU. Every rank has one copy stream per peer and phase, so the n-1 transfers of a chunk run side by side, and
// the whole round is replayed from one CUDA graph that spans all GPUs (one launch per round instead of hundreds of calls).
//
// Used by gpuprobe_allreduce_i32_ce (allreduce.cu, the whole buffer) and by the NVLS + copy-engine hybrid (nvls.cu, the
// tail of the buffer while the multimem kernel reduces the head).
#pragma once
#include <string>

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
        // Chunk boundaries inside a slice: equal chunks. GPUPROBE_CE_FIRST_PERMILLE gives the first chunk another share
        // (a bigger first chunk would hide the last add; measured, it loses: 2 ranks 740 / 742 / 713 / 676 / 654 GB/s for
        // 50 / 65 / 75 / 85 / 92 % -- profiles/r02_fabric.md).
        int first_permille = 1000 / chunks_;
        if (const char *e = getenv("GPUPROBE_CE_FIRST_PERMILLE")) first_permille = std::max(1, std::min(999, atoi(e)));
        bounds_.assign(chunks_ + 1, per_);
        bounds_[0] = 0;
        if (chunks_ > 1) {
            bounds_[1] = getenv("GPUPROBE_CE_FIRST_PERMILLE") ? std::min<uint64_t>(per_, (per_ / 1000u * static_cast<uint64_t>(first_permille) + 31) & ~31ull)
                                                              : std::min<uint64_t>(per_, ((per_ + chunks_ - 1) / chunks_ + 31) & ~31ull);
            const uint64_t rest = per_ - bounds_[1];
            for (int k = 2; k < chunks_; ++k)
                bounds_[k] = std::min<uint64_t>(per_, bounds_[1] + ((rest / (chunks_ - 1) * static_cast<uint64_t>(k - 1) + 31) & ~31ull));
        }
        // copy streams per rank and phase: one per peer up to 4 ranks (peer j uses stream (j-1) mod ns_), a single one
        // beyond -- more concurrent copies than that are more than the DMA engines serve well (8 ranks, 1 GiB: 576 GB/s
        // with 1 stream per phase, 469 with 2, 428 with 3, 424 with 7; profiles/r02_fabric.md).
        ns_ = nd_ <= 4 ? nd_ - 1 : 1;
        if (const char *e = getenv("GPUPROBE_CE_STREAMS")) ns_ = std::max(1, std::min(nd_ - 1, atoi(e)));
        S_.resize(nd_);
        cudaError_t e;
        for (int r = 0; r < nd_; ++r) {
            if ((e = cudaSetDevice(R_[r].ord)) != cudaSuccess) return e;
            Side &s = S_[r];
            int lo = 0, hi = 0;
            if ((e = cudaDeviceGetStreamPriorityRange(&lo, &hi)) != cudaSuccess) return e;
            // one copy stream per peer and phase: the n-1 transfers of a chunk run side by side, so the start-up gap
            // of one copy is covered by the others (serial copies on one stream lose ~15 us per copy)
            s.scatter.assign(ns_, nullptr);
            s.gather.assign(ns_, nullptr);
            for (auto &st : s.scatter)
                if ((e = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking)) != cudaSuccess) return e;
            for (auto &st : s.gather)
                if ((e = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking)) != cudaSuccess) return e;
            // the add must not queue behind a long-running kernel of the caller on the same GPU
            if ((e = cudaStreamCreateWithPriority(&s.add, cudaStreamNonBlocking, hi)) != cudaSuccess) return e;
            if ((e = cudaEventCreateWithFlags(&s.go, cudaEventDisableTiming)) != cudaSuccess) return e;
            s.gathered.assign(nd_ - 1, nullptr);
            for (auto &ev : s.gathered)
                if ((e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)) != cudaSuccess) return e;
            s.scattered.assign(static_cast<size_t>(chunks_) * (nd_ - 1), nullptr);
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
    //
    // A round is some hundred stream operations (n(n-1) peer copies per chunk and phase, events between them), and a peer
    // copy costs the host tens of microseconds to enqueue: issued call by call the host, not the links, bounds the round
    // from 4 ranks up (measured: 4 chunks at 4 ranks 520 GB/s against 695 with 2). So the round is captured ONCE into a
    // CUDA graph spanning all GPUs (stream capture forked from one origin stream through event waits) and replayed with
    // one launch per round. If the capture is refused the round is issued directly, as before.
    cudaError_t issue_round() {
        if (!graph_tried_) {
            build_graph();
            if (getenv("GPUPROBE_CE_DEBUG"))
                fprintf(stderr, "gpuprobe: copy-engine two-shot, %d ranks, %d chunks: %s\n", nd_, chunks_,
                        graph_exec_ ? "replaying a captured multi-GPU graph" : ("issuing call by call (" + graph_note_ + ")").c_str());
        }
        if (!graph_exec_) {
            std::vector<cudaStream_t> mains(nd_);
            for (int r = 0; r < nd_; ++r) mains[r] = R_[r].main;
            return issue_direct(mains);
        }
        cudaError_t e;
        if ((e = cudaSetDevice(R_[0].ord)) != cudaSuccess) return e;
        for (int r = 0; r < nd_; ++r) {  // the replay starts after what every rank's main stream holds now ...
            if ((e = cudaSetDevice(R_[r].ord)) != cudaSuccess) return e;
            if ((e = cudaEventRecord(S_[r].ext_go, R_[r].main)) != cudaSuccess) return e;
        }
        if ((e = cudaSetDevice(R_[0].ord)) != cudaSuccess) return e;
        for (int r = 0; r < nd_; ++r)
            if ((e = cudaStreamWaitEvent(origin_, S_[r].ext_go, 0)) != cudaSuccess) return e;
        if ((e = cudaGraphLaunch(graph_exec_, origin_)) != cudaSuccess) return e;
        if ((e = cudaEventRecord(ext_done_, origin_)) != cudaSuccess) return e;
        for (int r = 0; r < nd_; ++r) {  // ... and every main stream waits for the whole round
            if ((e = cudaSetDevice(R_[r].ord)) != cudaSuccess) return e;
            if ((e = cudaStreamWaitEvent(R_[r].main, ext_done_, 0)) != cudaSuccess) return e;
        }
        return cudaSuccess;
    }

    bool uses_graph() const { return graph_exec_ != nullptr; }

private:
    // Captures one round into a graph: origin_ forks every rank's internal main stream, issue_direct() does the rest.
    void build_graph() {
        graph_tried_ = true;
        if (const char *g = getenv("GPUPROBE_CE_GRAPH"); g && *g == '0') return;
        if (cudaSetDevice(R_[0].ord) != cudaSuccess) return;
        bool ok = cudaStreamCreateWithFlags(&origin_, cudaStreamNonBlocking) == cudaSuccess &&
                  cudaEventCreateWithFlags(&ext_done_, cudaEventDisableTiming) == cudaSuccess &&
                  cudaEventCreateWithFlags(&fork_, cudaEventDisableTiming) == cudaSuccess;
        std::vector<cudaStream_t> mains(nd_, nullptr);
        for (int r = 0; r < nd_ && ok; ++r) {
            ok = cudaSetDevice(R_[r].ord) == cudaSuccess && cudaStreamCreateWithFlags(&S_[r].gmain, cudaStreamNonBlocking) == cudaSuccess &&
                 cudaEventCreateWithFlags(&S_[r].ext_go, cudaEventDisableTiming) == cudaSuccess &&
                 cudaEventCreateWithFlags(&S_[r].join, cudaEventDisableTiming) == cudaSuccess;
            mains[r] = S_[r].gmain;
        }
        if (!ok) {
            (void)cudaGetLastError();
            return;
        }
        cudaGraph_t graph = nullptr;
        cudaSetDevice(R_[0].ord);
        const char *step = "cudaStreamBeginCapture";
        cudaError_t e = cudaStreamBeginCapture(origin_, cudaStreamCaptureModeRelaxed);
        if (e != cudaSuccess) {
            graph_note_ = std::string(step) + ": " + cudaGetErrorString(e);
            (void)cudaGetLastError();
            return;
        }
        step = "fork";
        e = cudaEventRecord(fork_, origin_);
        for (int r = 0; r < nd_ && e == cudaSuccess; ++r) {
            if ((e = cudaSetDevice(R_[r].ord)) != cudaSuccess) break;
            e = cudaStreamWaitEvent(mains[r], fork_, 0);
        }
        if (e == cudaSuccess) {
            step = "capturing the round";
            e = issue_direct(mains);
        }
        if (e == cudaSuccess) step = "join";
        for (int r = 0; r < nd_ && e == cudaSuccess; ++r) {  // every forked stream joins the origin again
            if ((e = cudaSetDevice(R_[r].ord)) != cudaSuccess) break;
            if ((e = cudaEventRecord(S_[r].join, mains[r])) != cudaSuccess) break;
        }
        if (e == cudaSuccess) e = cudaSetDevice(R_[0].ord);
        for (int r = 0; r < nd_ && e == cudaSuccess; ++r) e = cudaStreamWaitEvent(origin_, S_[r].join, 0);
        cudaSetDevice(R_[0].ord);
        const cudaError_t e_end = cudaStreamEndCapture(origin_, &graph);
        if (e == cudaSuccess && e_end != cudaSuccess) {
            step = "cudaStreamEndCapture";
            e = e_end;
        }
        if (e == cudaSuccess && graph) {
            step = "cudaGraphInstantiate";
            e = cudaGraphInstantiate(&graph_exec_, graph, 0);
        }
        if (graph) cudaGraphDestroy(graph);
        if (e != cudaSuccess) {
            graph_exec_ = nullptr;
            graph_note_ = std::string(step) + ": " + cudaGetErrorString(e);
            (void)cudaGetLastError();
        }
    }

    cudaError_t issue_direct(const std::vector<cudaStream_t> &mains) {
        cudaError_t e;
        for (int r = 0; r < nd_; ++r) {
            if ((e = cudaSetDevice(R_[r].ord)) != cudaSuccess) return e;
            if ((e = cudaEventRecord(S_[r].go, mains[r])) != cudaSuccess) return e;
            for (auto &st : S_[r].scatter)
                if ((e = cudaStreamWaitEvent(st, S_[r].go, 0)) != cudaSuccess) return e;
            for (auto &st : S_[r].gather)
                if ((e = cudaStreamWaitEvent(st, S_[r].go, 0)) != cudaSuccess) return e;
            if ((e = cudaStreamWaitEvent(S_[r].add, S_[r].go, 0)) != cudaSuccess) return e;
        }
        // stream j-1 of rank r serves peer (r + j) % n; event scattered[k*(n-1) + j-1] = "r's chunk k has reached that peer"
        for (int k = 0; k < chunks_; ++k)
            for (int r = 0; r < nd_; ++r) {
                if ((e = cudaSetDevice(R_[r].ord)) != cudaSuccess) return e;
                for (int j = 1; j < nd_; ++j) {
                    const int q = (r + j) % nd_;
                    uint64_t off, len;
                    chunk_range(slice_len(q), k, &off, &len);
                    if (len) {
                        const int slot = r < q ? r : r - 1;
                        // unified addressing tells the driver which two devices these are; this form is also what stream capture takes
                        e = cudaMemcpyAsync(R_[q].stage + static_cast<uint64_t>(slot) * per_ + off, R_[r].buf + slice_lo(q) + off,
                                            len * sizeof(uint32_t), cudaMemcpyDeviceToDevice, S_[r].scatter[(j - 1) % ns_]);
                        if (e != cudaSuccess) return e;
                    }
                    if ((e = cudaEventRecord(S_[r].scattered[static_cast<size_t>(k) * (nd_ - 1) + (j - 1)], S_[r].scatter[(j - 1) % ns_])) != cudaSuccess) return e;
                }
            }
        for (int k = 0; k < chunks_; ++k)  // add once every peer's copy of the chunk is here, then gather it to everyone
            for (int q = 0; q < nd_; ++q) {
                if ((e = cudaSetDevice(R_[q].ord)) != cudaSuccess) return e;
                for (int r = 0; r < nd_; ++r) {
                    if (r == q) continue;
                    const int j = (q - r + nd_) % nd_;  // q is peer j of rank r
                    if ((e = cudaStreamWaitEvent(S_[q].add, S_[r].scattered[static_cast<size_t>(k) * (nd_ - 1) + (j - 1)], 0)) != cudaSuccess) return e;
                }
                uint64_t off, len;
                chunk_range(slice_len(q), k, &off, &len);
                if (len) {
                    const uint64_t want = (len / 4 + 255) / 256;
                    const int grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(want, static_cast<uint64_t>(R_[q].sm_count) * 8)));
                    ce_add_staged_kernel<<<grid, 256, 0, S_[q].add>>>(R_[q].buf + slice_lo(q) + off, R_[q].stage + off, per_, nd_ - 1, len);
                    if ((e = cudaGetLastError()) != cudaSuccess) return e;
                }
                if ((e = cudaEventRecord(S_[q].added[k], S_[q].add)) != cudaSuccess) return e;
                for (int j = 1; j < nd_; ++j) {
                    const int p = (q + j) % nd_;
                    if ((e = cudaStreamWaitEvent(S_[q].gather[(j - 1) % ns_], S_[q].added[k], 0)) != cudaSuccess) return e;
                    if (!len) continue;
                    e = cudaMemcpyAsync(R_[p].buf + slice_lo(q) + off, R_[q].buf + slice_lo(q) + off, len * sizeof(uint32_t),
                                        cudaMemcpyDeviceToDevice, S_[q].gather[(j - 1) % ns_]);
                    if (e != cudaSuccess) return e;
                }
            }
        for (int r = 0; r < nd_; ++r) {
            if ((e = cudaSetDevice(R_[r].ord)) != cudaSuccess) return e;
            for (int j = 1; j < nd_; ++j)
                if ((e = cudaEventRecord(S_[r].gathered[j - 1], S_[r].gather[(j - 1) % ns_])) != cudaSuccess) return e;
        }
        for (int r = 0; r < nd_; ++r) {  // a rank's round is over when its outgoing gathers have left and the incoming ones have landed
            if ((e = cudaSetDevice(R_[r].ord)) != cudaSuccess) return e;
            for (int j = 1; j < nd_; ++j) {
                const int q = (r - j + nd_) % nd_;  // r is peer j of rank q
                if ((e = cudaStreamWaitEvent(mains[r], S_[r].gathered[j - 1], 0)) != cudaSuccess) return e;
                if ((e = cudaStreamWaitEvent(mains[r], S_[q].gathered[j - 1], 0)) != cudaSuccess) return e;
            }
        }
        return cudaSuccess;
    }

public:
    // The caller has synchronised the devices.
    void destroy() {
        if (graph_exec_) cudaGraphExecDestroy(graph_exec_);
        graph_exec_ = nullptr;
        if (!R_.empty() && cudaSetDevice(R_[0].ord) == cudaSuccess) {
            if (origin_) cudaStreamDestroy(origin_);
            if (ext_done_) cudaEventDestroy(ext_done_);
            if (fork_) cudaEventDestroy(fork_);
        }
        origin_ = nullptr;
        ext_done_ = fork_ = nullptr;
        for (int r = 0; r < static_cast<int>(S_.size()); ++r) {
            if (cudaSetDevice(R_[r].ord) != cudaSuccess) continue;
            Side &s = S_[r];
            for (cudaStream_t st : s.scatter)
                if (st) cudaStreamDestroy(st);
            for (cudaStream_t st : s.gather)
                if (st) cudaStreamDestroy(st);
            if (s.add) cudaStreamDestroy(s.add);
            if (s.go) cudaEventDestroy(s.go);
            for (cudaEvent_t ev : s.gathered)
                if (ev) cudaEventDestroy(ev);
            if (s.gmain) cudaStreamDestroy(s.gmain);
            if (s.ext_go) cudaEventDestroy(s.ext_go);
            if (s.join) cudaEventDestroy(s.join);
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
        std::vector<cudaStream_t> scatter, gather;  // ns_ each: stream (j-1) % ns_ serves rank (r + j) % n
        cudaStream_t add = nullptr;
        cudaStream_t gmain = nullptr;                  // this rank's main stream inside the captured graph
        cudaEvent_t ext_go = nullptr, join = nullptr;  // caller's main -> origin; gmain -> origin
        cudaEvent_t go = nullptr;
        std::vector<cudaEvent_t> scattered;  // [chunk * (n-1) + j-1]
        std::vector<cudaEvent_t> added;      // per chunk
        std::vector<cudaEvent_t> gathered;   // per peer: the last gather to that peer
    };
    uint64_t slice_lo(int q) const { return std::min<uint64_t>(per_ * static_cast<uint64_t>(q), count_); }
    uint64_t slice_len(int q) const { return std::min<uint64_t>(slice_lo(q) + per_, count_) - slice_lo(q); }
    // words [off, off+len) of chunk k inside a slice of `len_slice` words
    void chunk_range(uint64_t len_slice, int k, uint64_t *off, uint64_t *len) const {
        *off = std::min<uint64_t>(bounds_[k], len_slice);
        *len = std::min<uint64_t>(bounds_[k + 1], len_slice) - *off;
    }
    std::vector<CeRank> R_;
    std::vector<Side> S_;
    bool graph_tried_ = false;
    std::string graph_note_ = "graph not requested";
    cudaGraphExec_t graph_exec_ = nullptr;
    cudaStream_t origin_ = nullptr;
    cudaEvent_t ext_done_ = nullptr, fork_ = nullptr;
    int nd_ = 0, chunks_ = 2, ns_ = 1;
    uint64_t count_ = 0, per_ = 0;
    std::vector<uint64_t> bounds_;  // chunk k of a slice = words [bounds_[k], bounds_[k+1]) of it (multiples of 32)
};

}  // namespace gp
