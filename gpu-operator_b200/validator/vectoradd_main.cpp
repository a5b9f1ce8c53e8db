// vectoradd_main.cpp -- drop-in for the `vectorAdd` binary the validator's workload pods run
// (reference validator/cuda-workload-validation.yaml:19-20, validator/plugin-workload-validation.yaml:19-20,
// installed at validator/Dockerfile:49). Same contract: no arguments, exit 0 iff the check passes,
// stdout is free-form. A thin main over libgpuprobe:
//
//   always   the sample's own sequence at its own size: 50,000 fp32 elements generated on the host,
//            copied to device 0, added by the sm_100a kernel, copied back and compared with the host
//            loop h_A[i] + h_B[i] -- bit for bit (fp32 a+b rounds once), which subsumes the sample's
//            fabs(a+b-c) > 1e-5 test.
//   on request (env, all default off -- INTEGRATION.md section 4)
//            GPUPROBE_VADD_ELEMS   HBM bandwidth run on every visible GPU
//            GPUPROBE_P2P_BYTES    all-pairs NVLink copy+checksum (>= 2 GPUs)
//            GPUPROBE_AR_COUNT     int32 allreduce over all visible GPUs; GPUPROBE_AR_IMPL = nccl (default) | auto |
//                                  nvls | p2p | hybrid | push | ce | both.  If NCCL cannot be loaded (the validator image
//                                  strips the CUDA user libraries) the library's own kernel runs instead and the
//                                  result says so, unless GPUPROBE_AR_REQUIRE_NCCL=1 makes that a failure.
//            GPUPROBE_P2P_MODE     serial (default) | serial_push | rounds | rounds_push | serial_ce | rounds_ce
//            GPUPROBE_MIN_HBM_FRAC / GPUPROBE_MIN_LINK_FRAC  fail below this fraction of 8000 / 900 GB/s
//   Results go to $OUTPUT_DIR/gpuprobe-results.json and .prom when that directory exists.
#include <sys/stat.h>

#include <algorithm>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "../../include/gpuprobe.h"

static unsigned long long envU64(const char *name) {
    const char *v = getenv(name);
    return v && *v ? strtoull(v, nullptr, 0) : 0ull;
}
// JSON string body: NODE_NAME comes from the pod spec and is not ours to trust.
static std::string jsonEscape(const char *s) {
    std::string out;
    for (const unsigned char *p = reinterpret_cast<const unsigned char *>(s); *p; ++p) {
        char buf[8];
        if (*p == '"' || *p == '\\') {
            out += '\\';
            out += static_cast<char>(*p);
        } else if (*p < 0x20) {
            snprintf(buf, sizeof(buf), "\\u%04x", *p);
            out += buf;
        } else {
            out += static_cast<char>(*p);
        }
    }
    return out;
}
static double envF(const char *name) {
    const char *v = getenv(name);
    return v && *v ? atof(v) : 0.0;
}

int main(void) {
    char err[512] = "";
    int failed = 0;
    std::ostringstream json, prom;
    const char *node = getenv("NODE_NAME") ? getenv("NODE_NAME") : "";
    json << "{\"version\":2,\"node\":\"" << jsonEscape(node) << "\",\"timestamp\":" << static_cast<long long>(time(nullptr));
    if (gpuprobe_abi_version() != GPUPROBE_ABI_VERSION) {
        fprintf(stderr, "libgpuprobe ABI %d, this binary was built against %d\n", gpuprobe_abi_version(), GPUPROBE_ABI_VERSION);
        return EXIT_FAILURE;
    }

    // ---- the sample's check ----------------------------------------------------------------
    const int numElements = 50000;
    printf("[Vector addition of %d elements]\n", numElements);
    int ndev = 0;
    int rc = gpuprobe_device_count(&ndev);
    if (rc != GPUPROBE_OK || ndev < 1) {
        fprintf(stderr, "Failed to find a CUDA device (%s)!\n", gpuprobe_strerror(rc ? rc : GPUPROBE_ENODEV));
        return EXIT_FAILURE;
    }
    std::vector<float> h_A(numElements), h_B(numElements), h_C(numElements, -1.0f);
    gpuprobe_fill_inputs(h_A.data(), h_B.data(), 0, numElements, 0, GPUPROBE_GEN_UNIT);
    printf("Copy input data from the host memory to the CUDA device\n");
    rc = gpuprobe_vectoradd_buffers(0, h_A.data(), h_B.data(), h_C.data(), numElements, err, sizeof(err));
    if (rc != GPUPROBE_OK) {
        fprintf(stderr, "Failed to launch vectorAdd kernel (%s: %s)!\n", gpuprobe_strerror(rc), err);
        return EXIT_FAILURE;
    }
    printf("Copy output data from the CUDA device to the host memory\n");
    long long bad = 0;
    for (int i = 0; i < numElements; ++i) {
        const float want = h_A[i] + h_B[i];  // the host loop
        if (memcmp(&want, &h_C[i], sizeof(float)) != 0) {
            if (!bad) fprintf(stderr, "Result verification failed at element %d!\n", i);
            ++bad;
        }
    }
    json << ",\"vectoradd_check\":{\"n\":" << numElements << ",\"bit_exact\":" << (bad ? "false" : "true") << "}";
    if (bad) return EXIT_FAILURE;
    printf("Test PASSED\n");

    // ---- optional bandwidth / fabric probes ----------------------------------------------------
    const unsigned long long vaddElems = envU64("GPUPROBE_VADD_ELEMS");
    const double minHbm = envF("GPUPROBE_MIN_HBM_FRAC"), minLink = envF("GPUPROBE_MIN_LINK_FRAC");
    prom << "# HELP gpu_operator_node_cuda_probe_pass 1 if every requested gpuprobe check passed on the local node, 0 otherwise\n"
         << "# TYPE gpu_operator_node_cuda_probe_pass gauge\n";
    std::ostringstream promBody;
    if (vaddElems) {
        json << ",\"vectoradd\":[";
        promBody << "# HELP gpu_operator_node_cuda_hbm_gbps achieved vectorAdd HBM bandwidth (12 B/element) per GPU, GB/s\n"
                 << "# TYPE gpu_operator_node_cuda_hbm_gbps gauge\n";
        std::ostringstream frac;
        frac << "# HELP gpu_operator_node_cuda_hbm_roofline_fraction achieved vectorAdd bandwidth as a fraction of 8000 GB/s\n"
             << "# TYPE gpu_operator_node_cuda_hbm_roofline_fraction gauge\n";
        // every GPU at once, one host thread each (the library serialises per device, not across devices):
        // all HBM stacks of the node are loaded together, and the pass takes the time of one probe, not of ndev
        std::vector<gpuprobe_vadd_result> results(ndev);
        std::vector<int> rcs(ndev, GPUPROBE_OK);
        std::vector<std::string> errs(ndev);
        {
            std::vector<std::thread> workers;
            for (int d = 0; d < ndev; ++d)
                workers.emplace_back([&, d]() {
                    gpuprobe_vadd_cfg cfg{};
                    cfg.dev = d;
                    cfg.n_elems = vaddElems;
                    cfg.iters = 20;
                    cfg.warmup = 3;
                    cfg.verify = 3;
                    char e[512] = "";
                    rcs[d] = gpuprobe_vectoradd(&cfg, &results[d], e, sizeof(e));
                    errs[d] = e;
                });
            for (auto &w : workers) w.join();
        }
        for (int d = 0; d < ndev; ++d) {
            const gpuprobe_vadd_result &r = results[d];
            rc = rcs[d];
            const bool slow = rc == GPUPROBE_OK && minHbm > 0 && r.frac_nominal < minHbm && !r.l2_resident;
            if (rc != GPUPROBE_OK || slow) {
                ++failed;
                fprintf(stderr, "vectorAdd bandwidth probe on GPU %d failed: %s %s\n", d,
                        slow ? gpuprobe_strerror(GPUPROBE_FAIL_BANDWIDTH) : gpuprobe_strerror(rc), errs[d].c_str());
            }
            printf("GPU %d: vectorAdd n=%llu  %.1f GB/s (%.3f of 8000 GB/s nominal), %llu mismatches\n", d,
                   (unsigned long long)r.n, r.gbps_median, r.frac_nominal, (unsigned long long)r.mismatches);
            json << (d ? "," : "") << "{\"gpu\":" << d << ",\"n\":" << r.n << ",\"rc\":" << rc << ",\"gbps\":" << r.gbps_median
                 << ",\"frac_nominal\":" << r.frac_nominal << ",\"mismatches\":" << r.mismatches << ",\"sum64\":" << r.sum64
                 << ",\"l2_resident\":" << r.l2_resident << "}";
            promBody << "gpu_operator_node_cuda_hbm_gbps{node=\"@NODE@\",gpu=\"" << d << "\"} " << r.gbps_median << "\n";
            frac << "gpu_operator_node_cuda_hbm_roofline_fraction{node=\"@NODE@\",gpu=\"" << d << "\"} " << r.frac_nominal << "\n";
        }
        json << "]";
        promBody << frac.str();
    }
    const unsigned long long p2pBytes = envU64("GPUPROBE_P2P_BYTES");
    if (p2pBytes && ndev >= 2) {
        gpuprobe_p2p_cfg cfg{};
        cfg.ndev = ndev;
        cfg.bytes_per_pair = p2pBytes & ~15ull;
        cfg.iters = 10;
        cfg.warmup = 3;
        cfg.mode = GPUPROBE_P2P_SERIAL;
        if (const char *m = getenv("GPUPROBE_P2P_MODE")) {
            if (!strcmp(m, "serial_push")) cfg.mode = GPUPROBE_P2P_SERIAL_PUSH;
            else if (!strcmp(m, "rounds")) cfg.mode = GPUPROBE_P2P_ROUNDS;
            else if (!strcmp(m, "rounds_push")) cfg.mode = GPUPROBE_P2P_ROUNDS_PUSH;
            else if (!strcmp(m, "serial_ce")) cfg.mode = GPUPROBE_P2P_SERIAL_CE;
            else if (!strcmp(m, "rounds_ce")) cfg.mode = GPUPROBE_P2P_ROUNDS_CE;
        }
        std::vector<gpuprobe_p2p_cell> cells(static_cast<size_t>(ndev) * ndev);
        rc = gpuprobe_p2p_matrix(&cfg, cells.data(), err, sizeof(err));
        std::vector<double> g;
        bool allOk = true;
        for (int s = 0; s < ndev; ++s)
            for (int t = 0; t < ndev; ++t)
                if (s != t) {
                    g.push_back(cells[s * ndev + t].gbps);
                    allOk = allOk && cells[s * ndev + t].verified;
                }
        std::sort(g.begin(), g.end());
        const double mn = g.empty() ? 0 : g.front(), med = g.empty() ? 0 : g[g.size() / 2];
        const bool slow = rc == GPUPROBE_OK && minLink > 0 && mn / 900.0 < minLink;
        if (rc != GPUPROBE_OK || slow) {
            ++failed;
            fprintf(stderr, "P2P probe failed: %s %s\n", slow ? gpuprobe_strerror(GPUPROBE_FAIL_BANDWIDTH) : gpuprobe_strerror(rc), err);
        }
        printf("P2P: %zu directed pairs, min %.1f / median %.1f GB/s, checksums %s\n", g.size(), mn, med, allOk ? "ok" : "BAD");
        json << ",\"p2p\":{\"rc\":" << rc << ",\"pairs\":" << g.size() << ",\"bytes_per_pair\":" << cfg.bytes_per_pair
             << ",\"min_gbps\":" << mn << ",\"median_gbps\":" << med << ",\"all_verified\":" << (allOk ? "true" : "false") << "}";
        promBody << "# HELP gpu_operator_node_cuda_p2p_min_gbps slowest directed NVLink pair of the all-pairs copy+checksum probe, GB/s\n"
                 << "# TYPE gpu_operator_node_cuda_p2p_min_gbps gauge\n"
                 << "gpu_operator_node_cuda_p2p_min_gbps{node=\"@NODE@\"} " << mn << "\n";
    }
    if (p2pBytes && ndev == 1) {
        // one visible GPU (the plugin workload's pod): no link to probe, but the copy+checksum kernel itself can still be
        // checked, looped back on the device
        gpuprobe_p2p_cell cell{};
        rc = gpuprobe_p2p_loopback(0, p2pBytes & ~15ull, 3, 0, 0, &cell, err, sizeof(err));
        if (rc != GPUPROBE_OK) {
            ++failed;
            fprintf(stderr, "P2P kernel loopback failed: %s %s\n", gpuprobe_strerror(rc), err);
        }
        printf("P2P kernel loopback on the one visible GPU: %.1f GB/s local copy, checksum %s\n", cell.gbps, cell.verified ? "ok" : "BAD");
        json << ",\"p2p_loopback\":{\"rc\":" << rc << ",\"bytes\":" << (p2pBytes & ~15ull) << ",\"gbps\":" << cell.gbps
             << ",\"verified\":" << (cell.verified ? "true" : "false") << "}";
    }
    const unsigned long long arCount = envU64("GPUPROBE_AR_COUNT");
    if (arCount && ndev >= 2) {
        gpuprobe_ar_cfg cfg{};
        cfg.ndev = ndev;
        cfg.count = arCount;
        cfg.iters = 20;
        cfg.warmup = 5;
        gpuprobe_ar_result r{};
        // GPUPROBE_AR_IMPL: "nccl" (default; validates the library path), "auto" (the library's choice of hand-written
        // kernel for this rank count; "kernel" is the round-1 spelling), "nvls" / "p2p" / "hybrid" / "push" / "ce" (one of them),
        // "both" (NCCL result reported, NCCL and the auto kernel must both verify)
        const char *impl = getenv("GPUPROBE_AR_IMPL") ? getenv("GPUPROBE_AR_IMPL") : "nccl";
        const char *implRan = impl;
        if (!strcmp(impl, "nvls")) {
            rc = gpuprobe_allreduce_i32_nvls(&cfg, &r, err, sizeof(err));
        } else if (!strcmp(impl, "p2p")) {
            rc = gpuprobe_allreduce_i32_p2p(&cfg, &r, err, sizeof(err));
        } else if (!strcmp(impl, "hybrid")) {
            rc = gpuprobe_allreduce_i32_hybrid(&cfg, &r, err, sizeof(err));
        } else if (!strcmp(impl, "push")) {
            rc = gpuprobe_allreduce_i32_push(&cfg, &r, err, sizeof(err));
        } else if (!strcmp(impl, "ce") || !strcmp(impl, "copy_engine")) {
            rc = gpuprobe_allreduce_i32_ce(&cfg, &r, err, sizeof(err));
        } else if (!strcmp(impl, "auto") || !strcmp(impl, "kernel")) {
            rc = gpuprobe_allreduce_i32_auto(&cfg, &r, err, sizeof(err));
        } else {
            rc = gpuprobe_allreduce_i32(&cfg, &r, err, sizeof(err));
            if (rc == GPUPROBE_ENCCL && envU64("GPUPROBE_AR_REQUIRE_NCCL") == 0) {
                // No NCCL in this image: the fabric still gets its check, by the library's own kernel.
                fprintf(stderr, "NCCL unavailable (%s); running the library's allreduce kernel instead\n", err);
                json << ",\"allreduce_nccl\":\"unavailable\"";
                implRan = "auto (NCCL unavailable)";
                rc = gpuprobe_allreduce_i32_auto(&cfg, &r, err, sizeof(err));
            } else if (rc == GPUPROBE_OK && !strcmp(impl, "both")) {
                gpuprobe_ar_result r2{};
                rc = gpuprobe_allreduce_i32_auto(&cfg, &r2, err, sizeof(err));
                printf("allreduce int32 x%d (library kernel %d): busbw %.1f GB/s, %llu mismatches\n", r2.ndev, r2.impl, r2.busbw_gbps,
                       (unsigned long long)r2.mismatches);
                json << ",\"allreduce_kernel\":{\"rc\":" << rc << ",\"impl\":" << r2.impl << ",\"busbw_gbps\":" << r2.busbw_gbps
                     << ",\"mismatches\":" << r2.mismatches << ",\"sum64\":" << r2.sum64 << "}";
                // both paths reduced the same inputs for the same number of rounds: their outputs must be the same bits
                if (rc == GPUPROBE_OK && r2.rounds == r.rounds && (r2.sum64 != r.sum64 || r2.xor32 != r.xor32)) {
                    snprintf(err, sizeof(err), "NCCL and the library kernel disagree: %016llx vs %016llx",
                             (unsigned long long)r.sum64, (unsigned long long)r2.sum64);
                    rc = GPUPROBE_FAIL_VERIFY;
                }
            }
        }
        const bool slow = rc == GPUPROBE_OK && minLink > 0 && r.frac_of_peak < minLink;
        if (rc != GPUPROBE_OK || slow) {
            ++failed;
            fprintf(stderr, "allreduce probe failed: %s %s\n", slow ? gpuprobe_strerror(GPUPROBE_FAIL_BANDWIDTH) : gpuprobe_strerror(rc), err);
        }
        printf("allreduce int32 x%d (%s): busbw %.1f GB/s, %llu mismatches\n", r.ndev, implRan, r.busbw_gbps, (unsigned long long)r.mismatches);
        json << ",\"allreduce\":{\"rc\":" << rc << ",\"impl\":" << r.impl << ",\"ndev\":" << r.ndev << ",\"bytes\":" << r.bytes
             << ",\"busbw_gbps\":" << r.busbw_gbps << ",\"mismatches\":" << r.mismatches << ",\"rounds\":" << r.rounds
             << ",\"sum64\":" << r.sum64 << ",\"xor32\":" << r.xor32 << ",\"ranks_equal\":" << r.ranks_equal << "}";
        promBody << "# HELP gpu_operator_node_cuda_allreduce_busbw_gbps int32 allreduce bus bandwidth over all GPUs of the node, GB/s\n"
                 << "# TYPE gpu_operator_node_cuda_allreduce_busbw_gbps gauge\n"
                 << "gpu_operator_node_cuda_allreduce_busbw_gbps{node=\"@NODE@\"} " << r.busbw_gbps << "\n";
    }
    json << ",\"pass\":" << (failed ? "false" : "true") << "}\n";
    prom << "gpu_operator_node_cuda_probe_pass{node=\"@NODE@\"} " << (failed ? 0 : 1) << "\n" << promBody.str();

    const char *outDir = getenv("OUTPUT_DIR") ? getenv("OUTPUT_DIR") : "/run/nvidia/validations";
    struct stat st;
    if (stat(outDir, &st) == 0 && S_ISDIR(st.st_mode)) {
        for (int k = 0; k < 2; ++k) {
            const char *base = getenv("GPUPROBE_RESULTS_BASENAME");  // set for the plugin workload (one GPU)
            std::string path = std::string(outDir) + "/" + (base && *base ? base : "gpuprobe-results") + (k ? ".prom" : ".json");
            std::string tmp = path + ".tmp";
            if (FILE *fp = fopen(tmp.c_str(), "w")) {
                const std::string body = k ? prom.str() : json.str();
                fwrite(body.data(), 1, body.size(), fp);
                fclose(fp);
                rename(tmp.c_str(), path.c_str());
            } else {
                fprintf(stderr, "warning: cannot write %s: %s\n", tmp.c_str(), strerror(errno));
            }
        }
    } else if (vaddElems || p2pBytes || arCount) {
        // In a cluster this means the workload pod does not mount the validations hostPath (integration/patches/0006):
        // the probes ran, but the metrics exporter will never see their figures.
        fprintf(stderr, "warning: %s is not a directory; probe results were not written\n", outDir);
    }
    gpuprobe_shutdown();
    if (failed) return EXIT_FAILURE;
    printf("Done\n");
    return EXIT_SUCCESS;
}
