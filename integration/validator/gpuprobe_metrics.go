// validator/gpuprobe_metrics.go -- gauges for the probe's achieved bandwidth, in the style of
// validator/metrics.go (promauto gauge vectors labelled {node}, a watcher goroutine per source).
//
// NOT COMPILED (no Go toolchain here); the C++ mirror's metrics component serves the same gauge
// names and is tested in tests/test_validator_host.py.

package main

import (
	"encoding/json"
	"os"
	"path/filepath"
	"strconv"
	"time"

	promcli "github.com/prometheus/client_golang/prometheus"
	"github.com/prometheus/client_golang/prometheus/promauto"
	log "github.com/sirupsen/logrus"
)

const probeResultsCheckDelaySeconds = 30

var (
	probePass = promauto.NewGaugeVec(promcli.GaugeOpts{
		Name: "gpu_operator_node_cuda_probe_pass",
		Help: "1 if every requested gpuprobe check passed on the local node, 0 otherwise",
	}, []string{"node"})
	probeHBM = promauto.NewGaugeVec(promcli.GaugeOpts{
		Name: "gpu_operator_node_cuda_hbm_gbps",
		Help: "achieved vectorAdd HBM bandwidth (12 B/element) per GPU, GB/s",
	}, []string{"node", "gpu"})
	probeHBMFrac = promauto.NewGaugeVec(promcli.GaugeOpts{
		Name: "gpu_operator_node_cuda_hbm_roofline_fraction",
		Help: "achieved vectorAdd bandwidth as a fraction of 8000 GB/s",
	}, []string{"node", "gpu"})
	probeP2PMin = promauto.NewGaugeVec(promcli.GaugeOpts{
		Name: "gpu_operator_node_cuda_p2p_min_gbps",
		Help: "slowest directed NVLink pair of the all-pairs copy+checksum probe, GB/s",
	}, []string{"node"})
	probeARBusBW = promauto.NewGaugeVec(promcli.GaugeOpts{
		Name: "gpu_operator_node_cuda_allreduce_busbw_gbps",
		Help: "int32 allreduce bus bandwidth over all GPUs of the node, GB/s",
	}, []string{"node"})
)

// watchProbeResults re-reads <OUTPUT_DIR>/gpuprobe-results.json and republishes it. The file is written by the cuda
// workload pod; it reaches this pod because both mount the node's /run/nvidia/validations -- the exporter through
// /run/nvidia (assets/state-node-status-exporter/0700_daemonset.yaml:49-51), the workload pod through the hostPath
// that forwardProbeEnv / integration/patches/0006 add.
func (nm *NodeMetrics) watchProbeResults() {
	path := filepath.Join(outputDirFlag, probeResultsFile)
	log.Printf("metrics: gpuprobe: watching %s", path)
	for {
		if data, err := os.ReadFile(path); err == nil {
			var res ProbeResults
			if json.Unmarshal(data, &res) == nil {
				pass := 0.0
				if res.Pass {
					pass = 1.0
				}
				probePass.WithLabelValues(nodeNameFlag).Set(pass)
				for _, v := range res.VectorAdd {
					gpu := strconv.Itoa(v.GPU)
					probeHBM.WithLabelValues(nodeNameFlag, gpu).Set(v.GBps)
					probeHBMFrac.WithLabelValues(nodeNameFlag, gpu).Set(v.FracNominal)
				}
				if res.P2P != nil {
					probeP2PMin.WithLabelValues(nodeNameFlag).Set(res.P2P.MinGBps)
				}
				if res.AllReduce != nil {
					probeARBusBW.WithLabelValues(nodeNameFlag).Set(res.AllReduce.BusBWGBps)
				}
			}
		}
		time.Sleep(probeResultsCheckDelaySeconds * time.Second)
	}
}
