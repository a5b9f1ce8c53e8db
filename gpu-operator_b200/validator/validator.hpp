// validator.hpp -- C++ host-side mirror of the reference validator's interface for the one
// path this repo rebuilds (reference validator/main.go, validator/metrics.go).
//
// The reference is Go and there is no Go toolchain on either box, so the host side above the
// C ABI is written in C++ with the same names, argument meaning and error behaviour: the same
// flags and env fallbacks (main.go:212-315), the same validateFlags messages (:339-370), the
// same status files (:123-157, :710-741), the same component flow for `cuda` (:1176-1293) and
// `plugin` (:803-845, :931-1015, :1083-1161), the same retry budgets (:159-165, :1042-1059).
// What is NOT mirrored is the Kubernetes API client (control plane, out of scope): the two
// things the validator asks the API server for are behind small interfaces --
//   WorkloadLauncher  "create this pod and tell me when its phase is Succeeded"
//   NodeResources     "what does node.Status.Capacity / Allocatable hold"
// with headless implementations (a child process; the local device count or a capacity file).
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace validator {

// ---- constants (reference validator/main.go:123-205) ------------------------------------
constexpr const char *defaultStatusPath = "/run/nvidia/validations";
constexpr int defaultSleepIntervalSeconds = 5;
constexpr int defaultMetricsPort = 0;
constexpr const char *driverStatusFile = "driver-ready";
constexpr const char *hostDriverStatusFile = "host-driver-ready";
constexpr const char *toolkitStatusFile = "toolkit-ready";
constexpr const char *pluginStatusFile = "plugin-ready";
constexpr const char *cudaStatusFile = "cuda-ready";
constexpr int podCreationWaitRetries = 60;
constexpr int podCreationSleepIntervalSeconds = 5;
constexpr int gpuResourceDiscoveryWaitRetries = 30;
constexpr int gpuResourceDiscoveryIntervalSeconds = 5;
constexpr const char *genericGPUResourceType = "nvidia.com/gpu";
constexpr const char *migGPUResourcePrefix = "nvidia.com/mig-";
constexpr const char *pluginWorkloadPodSpecPath = "/var/nvidia/manifests/plugin-workload-validation.yaml";
constexpr const char *cudaWorkloadPodSpecPath = "/var/nvidia/manifests/cuda-workload-validation.yaml";
constexpr const char *cudaValidatorLabelValue = "nvidia-cuda-validator";
constexpr const char *pluginValidatorLabelValue = "nvidia-device-plugin-validator";
// new: where the probe leaves its numbers for the metrics component
constexpr const char *probeResultsJSON = "gpuprobe-results.json";
constexpr const char *probeResultsProm = "gpuprobe-results.prom";

// ---- flags (reference validator/main.go:107-121, 212-315) -------------------------------
struct Flags {
    std::string kubeconfig;
    std::string nodeName;
    std::string namespace_;
    bool withWait = false;
    bool withWorkload = false;
    std::string component;
    bool cleanupAll = false;
    std::string outputDir = defaultStatusPath;
    int sleepIntervalSeconds = defaultSleepIntervalSeconds;
    std::string migStrategy = "single";
    int metricsPort = defaultMetricsPort;
    std::string defaultGPUWorkloadConfig;
    bool disableDevCharSymlinkCreation = false;
};

// Parses argv with env fallbacks exactly as urfave/cli does for the reference's flag table.
// Returns "" or an error message.
std::string parseFlags(int argc, char **argv, Flags *f);
// reference validateFlags (main.go:339-370): "" or the reference's error text.
std::string validateFlags(const Flags &f);
bool isValidComponent(const std::string &c);  // main.go:372-397

// ---- status files (main.go:710-741) ------------------------------------------------------
std::string createStatusFile(const std::string &path);  // "" or "unable to create status file ..."
std::string deleteStatusFile(const std::string &path);  // "" or "unable to remove driver status file ..."

// ---- pod spec (main.go:1061-1081 loadPodSpec) --------------------------------------------
struct PodSpec {
    std::string appLabel;
    std::string generateName;
    std::string namespace_;
    std::string nodeName;
    std::string image;
    std::string imagePullPolicy;
    std::vector<std::string> imagePullSecrets;
    std::string runtimeClassName;
    std::string restartPolicy;
    std::string initName;
    std::vector<std::string> initCommand;  // e.g. ["sh","-c"]
    std::vector<std::string> initArgs;     // e.g. ["vectorAdd"]
    std::map<std::string, std::string> initEnv;
    std::map<std::string, std::string> initResourceLimits;  // plugin: {resourceName: "1"}
    std::map<std::string, std::string> initVolumeMounts;    // init container: volume name -> mountPath
    std::map<std::string, std::string> hostPathVolumes;     // spec.volumes: volume name -> hostPath.path
};
// New: the workload pod must mount the validations hostPath, or the probe's results file never leaves the pod (the
// reference's workload pods declare no volumes: validator/cuda-workload-validation.yaml:1-35). Adds the volume and the
// init container's mount when the manifest lacks them -- what the Go forwardProbeEnv does (integration/validator/gpuprobe.go).
void ensureValidationsVolume(PodSpec *pod);
// True if the init container mounts a hostPath volume at `dir`.
bool mountsHostDir(const PodSpec &pod, const std::string &dir);
// Reads the subset of the Pod manifest the validator touches. Like the reference it treats an
// unreadable file as fatal (the reference panics, main.go:1064-1066): returns an error string.
// This is a reader for the block-style YAML subset the two shipped manifests use, not a YAML parser: nested maps by
// indentation, "- " sequences of maps, flow or block lists for command/args, comments outside quotes. Anchors and
// aliases, merge keys, flow mappings, block scalars (| and >), tabs and a second document are rejected with an error
// that names the construct and the line.
std::string loadPodSpec(const std::string &path, PodSpec *pod);

// ---- the two seams to the Kubernetes API --------------------------------------------------
using ResourceList = std::map<std::string, int64_t>;  // v1.ResourceList with Quantity.Value()

class NodeResources {
public:
    virtual ~NodeResources() = default;
    // node.Status.Capacity / Allocatable of NODE_NAME; error text on failure
    virtual std::string get(ResourceList *capacity, ResourceList *allocatable) = 0;
};
// Headless: GPUPROBE_NODE_CAPACITY_FILE ("name quantity" per line) if set, else
// {"nvidia.com/gpu": <CUDA device count>} from libgpuprobe.
std::unique_ptr<NodeResources> makeLocalNodeResources();

class WorkloadLauncher {
public:
    virtual ~WorkloadLauncher() = default;
    // Delete a stale pod with this app label on this node, create the pod; returns its name.
    virtual std::string create(const PodSpec &pod, std::string *name) = 0;
    // "Succeeded", "Running", "Pending", "Failed" (pod.Status.Phase)
    virtual std::string phase(const std::string &name, std::string *phase) = 0;
};
// Headless: the pod's init container becomes a child process `sh -c <args>`; restartPolicy
// OnFailure restarts it after a failure, as the kubelet would.
std::unique_ptr<WorkloadLauncher> makeLocalLauncher(const Flags &f);

// waitForPod (main.go:1042-1059): polls the phase; pollMillis is 5000 in the reference.
std::string waitForPod(WorkloadLauncher &l, const std::string &name, int retries, int pollMillis);

// ---- Plugin resource logic (main.go:1083-1161) -------------------------------------------
std::string availableMIGResourceName(const ResourceList &r);
std::string availableGenericResourceName(const ResourceList &r);
int64_t countGPUResources(const ResourceList &capacity);
std::string getGPUResourceName(const ResourceList &allocatable, std::string *name);

// ---- components ---------------------------------------------------------------------------
struct Runtime {
    Flags flags;
    std::unique_ptr<WorkloadLauncher> launcher;
    std::unique_ptr<NodeResources> node;
    int pollMillis = podCreationSleepIntervalSeconds * 1000;
    int discoveryPollMillis = gpuResourceDiscoveryIntervalSeconds * 1000;
    // A shorter poll keeps the reference's time budgets (60 x 5 s, 30 x 5 s) by scaling the retry
    // counts, unless keepRetryCounts asks for the reference's counts verbatim (tests).
    bool keepRetryCounts = false;
    std::string manifestDir;  // "" -> the reference's absolute paths
};
std::string validateCUDA(Runtime &rt);     // CUDA.validate   main.go:1176-1211
std::string validatePlugin(Runtime &rt);   // Plugin.validate main.go:803-845
std::string validateDriver(Runtime &rt);   // Driver.validate main.go:637-691 (nvidia-smi check only)
std::string validateToolkit(Runtime &rt);  // Toolkit.validate main.go:775-801
std::string runMetrics(Runtime &rt);       // Metrics.run     main.go:1295-1299 + metrics.go
// start (main.go:439-545): cleanup-all, mkdir, dispatch. "" or "error validating ...: <cause>"
std::string start(Runtime &rt);

// one scrape of the node metrics in Prometheus text format (metrics.go gauges + probe gauges)
std::string renderMetrics(const Flags &f, double metricsReadyTs);

void logInfo(const char *fmt, ...);
void logError(const char *fmt, ...);

}  // namespace validator
