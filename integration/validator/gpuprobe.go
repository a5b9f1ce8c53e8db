// validator/gpuprobe.go -- cgo binding of libgpuprobe and the `cuda-workload` component.
//
// NOT COMPILED: neither the build container nor the GPU box of this repository has a Go
// toolchain. Every C call made here is exercised by the repository's tests through the ctypes
// mirror (gpu-operator_b200/gpuprobe.py), which binds the same symbols with the same struct
// layouts, and by the C++ host programs under gpu-operator_b200/validator/, whose vectorAdd
// main is the same sequence as runProbes below.
//
// Drop into the reference's validator/ directory together with the patches in
// integration/patches/ (see INTEGRATION.md).

package main

/*
#cgo CFLAGS: -I${SRCDIR}/../gpuprobe/include
#cgo LDFLAGS: -lgpuprobe
#include <stdlib.h>
#include "gpuprobe.h"
*/
import "C"

import (
	"encoding/json"
	"fmt"
	"os"
	"path/filepath"
	"sort"
	"strconv"
	"strings"
	"unsafe"

	log "github.com/sirupsen/logrus"
	v1 "k8s.io/api/core/v1"
)

const (
	probeEnvPrefix       = "GPUPROBE_"
	probeResultsFile     = "gpuprobe-results.json"
	referenceNumElements = 50000 // the size the cuda-samples vectorAdd binary uses
	nominalHBMGBps       = 8000.0
	nominalLinkGBps      = 900.0
)

const validationsVolumeName = "run-nvidia-validations"

// forwardProbeEnv copies every GPUPROBE_* variable of the validator's own environment into the
// workload pod's init container (validator.cuda.env / validator.plugin.env only reach the DaemonSet),
// and makes sure the pod can hand its results back: the workload manifests of the reference declare no
// volumes (validator/cuda-workload-validation.yaml), so without the hostPath below gpuprobe-results.json
// would stay inside the pod and the exporter's gauges would never get a value.
func forwardProbeEnv(pod *v1.Pod) {
	for _, kv := range os.Environ() {
		if !strings.HasPrefix(kv, probeEnvPrefix) {
			continue
		}
		parts := strings.SplitN(kv, "=", 2)
		pod.Spec.InitContainers[0].Env = append(pod.Spec.InitContainers[0].Env, v1.EnvVar{Name: parts[0], Value: parts[1]})
	}
	pod.Spec.InitContainers[0].Env = append(pod.Spec.InitContainers[0].Env,
		v1.EnvVar{Name: "OUTPUT_DIR", Value: outputDirFlag}, v1.EnvVar{Name: "NODE_NAME", Value: nodeNameFlag})
	ensureValidationsVolume(pod)
}

// ensureValidationsVolume adds the validations hostPath (the one the validator DaemonSet itself mounts,
// assets/state-operator-validation/0500_daemonset.yaml:50-51,158-162) and the init container's mount, unless the
// manifest already has them (integration/patches/0006 adds them to both workload manifests).
func ensureValidationsVolume(pod *v1.Pod) {
	for _, m := range pod.Spec.InitContainers[0].VolumeMounts {
		if m.MountPath == outputDirFlag {
			return
		}
	}
	hostPathType := v1.HostPathDirectoryOrCreate
	haveVolume := false
	for _, vol := range pod.Spec.Volumes {
		haveVolume = haveVolume || vol.Name == validationsVolumeName
	}
	if !haveVolume {
		pod.Spec.Volumes = append(pod.Spec.Volumes, v1.Volume{Name: validationsVolumeName,
			VolumeSource: v1.VolumeSource{HostPath: &v1.HostPathVolumeSource{Path: outputDirFlag, Type: &hostPathType}}})
	}
	pod.Spec.InitContainers[0].VolumeMounts = append(pod.Spec.InitContainers[0].VolumeMounts,
		v1.VolumeMount{Name: validationsVolumeName, MountPath: outputDirFlag})
}

func probeError(where string, rc C.int, msg *C.char) error {
	if rc == C.GPUPROBE_OK {
		return nil
	}
	return fmt.Errorf("%s: %s: %s", where, C.GoString(C.gpuprobe_strerror(rc)), C.GoString(msg))
}

func envUint(name string) uint64 {
	v, _ := strconv.ParseUint(os.Getenv(name), 0, 64)
	return v
}

func envFloat(name string) float64 {
	v, _ := strconv.ParseFloat(os.Getenv(name), 64)
	return v
}

// ProbeResults is what the metrics component reads back (validator/gpuprobe_metrics.go).
type ProbeResults struct {
	Version   int              `json:"version"`
	Node      string           `json:"node"`
	Pass      bool             `json:"pass"`
	VectorAdd []VectorAddEntry `json:"vectoradd,omitempty"`
	P2P       *P2PEntry        `json:"p2p,omitempty"`
	AllReduce *AllReduceEntry  `json:"allreduce,omitempty"`
}

type VectorAddEntry struct {
	GPU         int     `json:"gpu"`
	N           uint64  `json:"n"`
	GBps        float64 `json:"gbps"`
	FracNominal float64 `json:"frac_nominal"`
	Mismatches  uint64  `json:"mismatches"`
}

type P2PEntry struct {
	Pairs       int     `json:"pairs"`
	MinGBps     float64 `json:"min_gbps"`
	MedianGBps  float64 `json:"median_gbps"`
	AllVerified bool    `json:"all_verified"`
}

type AllReduceEntry struct {
	NDev       int     `json:"ndev"`
	Impl       int     `json:"impl"` // GPUPROBE_AR_IMPL_*
	BusBWGBps  float64 `json:"busbw_gbps"`
	Mismatches uint64  `json:"mismatches"`
	Rounds     uint32  `json:"rounds"`
	Sum64      uint64  `json:"sum64"` // digest of the reduced buffer (ABI 2)
	Xor32      uint32  `json:"xor32"`
	RanksEqual bool    `json:"ranks_equal"`
}

// vectorAddHostCheck is the sample's own sequence at its own size, checked by a Go host loop:
// fp32 a+b rounds once, so == on the results is the right test.
func vectorAddHostCheck(dev int, n int) error {
	a, b, c := make([]float32, n), make([]float32, n), make([]float32, n)
	C.gpuprobe_fill_inputs((*C.float)(&a[0]), (*C.float)(&b[0]), 0, C.uint64_t(n), 0, C.GPUPROBE_GEN_UNIT)
	msg := (*C.char)(C.calloc(512, 1))
	defer C.free(unsafe.Pointer(msg))
	rc := C.gpuprobe_vectoradd_buffers(C.int(dev), (*C.float)(&a[0]), (*C.float)(&b[0]), (*C.float)(&c[0]), C.uint64_t(n), msg, 512)
	if err := probeError("vectorAdd", rc, msg); err != nil {
		return err
	}
	for i := range c {
		if c[i] != a[i]+b[i] {
			return fmt.Errorf("vectorAdd: result verification failed at element %d: got %g want %g", i, c[i], a[i]+b[i])
		}
	}
	return nil
}

// runProbes is the `cuda-workload` component: what the workload pod's init container runs instead of `vectorAdd`.
func runProbes() error {
	var ndevC C.int
	if rc := C.gpuprobe_device_count(&ndevC); rc != C.GPUPROBE_OK || ndevC < 1 {
		return fmt.Errorf("failed to find a CUDA device: %s", C.GoString(C.gpuprobe_strerror(rc)))
	}
	ndev := int(ndevC)
	if v := int(C.gpuprobe_abi_version()); v != C.GPUPROBE_ABI_VERSION {
		return fmt.Errorf("libgpuprobe ABI %d, this binary was built against %d", v, int(C.GPUPROBE_ABI_VERSION))
	}
	if err := vectorAddHostCheck(0, referenceNumElements); err != nil {
		return err
	}
	log.Infof("vectorAdd of %d elements on device 0: bit-exact", referenceNumElements)

	res := ProbeResults{Version: 2, Node: nodeNameFlag, Pass: true}
	msg := (*C.char)(C.calloc(512, 1))
	defer C.free(unsafe.Pointer(msg))
	var failures []string
	minHBM, minLink := envFloat("GPUPROBE_MIN_HBM_FRAC"), envFloat("GPUPROBE_MIN_LINK_FRAC")

	if n := envUint("GPUPROBE_VADD_ELEMS"); n > 0 {
		for d := 0; d < ndev; d++ {
			cfg := C.gpuprobe_vadd_cfg{dev: C.int(d), n_elems: C.uint64_t(n), iters: 20, warmup: 3, verify: 3}
			var out C.gpuprobe_vadd_result
			rc := C.gpuprobe_vectoradd(&cfg, &out, msg, 512)
			if err := probeError(fmt.Sprintf("vectorAdd on GPU %d", d), rc, msg); err != nil {
				failures = append(failures, err.Error())
			} else if minHBM > 0 && out.l2_resident == 0 && float64(out.frac_nominal) < minHBM {
				failures = append(failures, fmt.Sprintf("GPU %d: %.1f GB/s is below %.2f of %.0f GB/s", d, float64(out.gbps_median), minHBM, nominalHBMGBps))
			}
			res.VectorAdd = append(res.VectorAdd, VectorAddEntry{GPU: d, N: uint64(out.n), GBps: float64(out.gbps_median),
				FracNominal: float64(out.frac_nominal), Mismatches: uint64(out.mismatches)})
		}
	}
	if bytes := envUint("GPUPROBE_P2P_BYTES"); bytes > 0 && ndev >= 2 {
		cfg := C.gpuprobe_p2p_cfg{ndev: C.int(ndev), bytes_per_pair: C.uint64_t(bytes &^ 15), iters: 10, warmup: 3, mode: C.GPUPROBE_P2P_SERIAL}
		cells := make([]C.gpuprobe_p2p_cell, ndev*ndev)
		rc := C.gpuprobe_p2p_matrix(&cfg, &cells[0], msg, 512)
		var g []float64
		ok := true
		for s := 0; s < ndev; s++ {
			for t := 0; t < ndev; t++ {
				if s != t {
					g = append(g, float64(cells[s*ndev+t].gbps))
					ok = ok && cells[s*ndev+t].verified != 0
				}
			}
		}
		sort.Float64s(g)
		if err := probeError("p2p matrix", rc, msg); err != nil {
			failures = append(failures, err.Error())
		} else if minLink > 0 && g[0]/nominalLinkGBps < minLink {
			failures = append(failures, fmt.Sprintf("slowest NVLink pair %.1f GB/s is below %.2f of %.0f GB/s", g[0], minLink, nominalLinkGBps))
		}
		res.P2P = &P2PEntry{Pairs: len(g), MinGBps: g[0], MedianGBps: g[len(g)/2], AllVerified: ok}
	}
	if count := envUint("GPUPROBE_AR_COUNT"); count > 0 && ndev >= 2 {
		cfg := C.gpuprobe_ar_cfg{ndev: C.int(ndev), count: C.uint64_t(count), iters: 20, warmup: 5}
		var out C.gpuprobe_ar_result
		rc := C.gpuprobe_allreduce_i32(&cfg, &out, msg, 512)
		if rc == C.GPUPROBE_ENCCL && os.Getenv("GPUPROBE_AR_REQUIRE_NCCL") != "1" {
			// the validator image ships no libnccl.so.2 unless it was built with GPUPROBE_NCCL_STATIC=1: the fabric
			// still gets its check, by the library's own allreduce kernel
			log.Warnf("NCCL unavailable (%s); running the library's allreduce kernel instead", C.GoString(msg))
			rc = C.gpuprobe_allreduce_i32_auto(&cfg, &out, msg, 512)
		}
		if err := probeError("allreduce", rc, msg); err != nil {
			failures = append(failures, err.Error())
		} else if minLink > 0 && float64(out.frac_of_peak) < minLink {
			failures = append(failures, fmt.Sprintf("allreduce busbw %.1f GB/s is below %.2f of %.0f GB/s", float64(out.busbw_gbps), minLink, nominalLinkGBps))
		}
		res.AllReduce = &AllReduceEntry{NDev: int(out.ndev), Impl: int(out.impl), BusBWGBps: float64(out.busbw_gbps),
			Mismatches: uint64(out.mismatches), Rounds: uint32(out.rounds), Sum64: uint64(out.sum64), Xor32: uint32(out.xor32),
			RanksEqual: out.ranks_equal != 0}
	}

	res.Pass = len(failures) == 0
	if data, err := json.Marshal(res); err == nil {
		tmp := filepath.Join(outputDirFlag, probeResultsFile+".tmp")
		if err := os.WriteFile(tmp, data, 0644); err != nil {
			// the probes ran, but nobody will see their figures: say so (a missing hostPath mount is the usual cause)
			log.Warnf("cannot write %s: %v", tmp, err)
		} else if err := os.Rename(tmp, filepath.Join(outputDirFlag, probeResultsFile)); err != nil {
			log.Warnf("cannot publish %s: %v", probeResultsFile, err)
		}
	}
	C.gpuprobe_shutdown()
	if !res.Pass {
		return fmt.Errorf("%s", strings.Join(failures, "; "))
	}
	return nil
}
