// validator.cpp -- see validator.hpp. Reference line numbers are into validator/main.go unless noted.
#include "validator.hpp"

#include <arpa/inet.h>
#include <dirent.h>
#include <netinet/in.h>
#include <signal.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

#include <cerrno>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <fstream>
#include <mutex>
#include <random>
#include <sstream>
#include <thread>

#include "../../include/gpuprobe.h"

extern char **environ;

namespace validator {

// ------------------------------------------------------------------------------ logging
static void vlog(const char *level, const char *fmt, va_list ap) {
    char msg[2048];
    vsnprintf(msg, sizeof(msg), fmt, ap);
    char ts[64];
    time_t now = time(nullptr);
    struct tm tmv;
    gmtime_r(&now, &tmv);
    strftime(ts, sizeof(ts), "%Y-%m-%dT%H:%M:%SZ", &tmv);
    fprintf(stderr, "time=\"%s\" level=%s msg=\"%s\"\n", ts, level, msg);  // logrus text format
}
void logInfo(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vlog("info", fmt, ap);
    va_end(ap);
}
void logError(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vlog("error", fmt, ap);
    va_end(ap);
}

static std::string sprintfStr(const char *fmt, ...) {
    char buf[4096];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    return buf;
}

static bool fileExists(const std::string &p) {
    struct stat st;
    return stat(p.c_str(), &st) == 0;
}

static const char *envOr(const char *name, const char *dflt) {
    const char *v = getenv(name);
    return v ? v : dflt;
}

// ------------------------------------------------------------------------------ flags
namespace {
struct FlagDef {
    const char *name;
    const char *alias;
    const char *env;
    enum Kind { STRING, BOOL, INT } kind;
};
const FlagDef kFlags[] = {
    {"kubeconfig", "", "KUBECONFIG", FlagDef::STRING},
    {"node-name", "n", "NODE_NAME", FlagDef::STRING},
    {"namespace", "ns", "OPERATOR_NAMESPACE", FlagDef::STRING},
    {"with-wait", "w", "WITH_WAIT", FlagDef::BOOL},
    {"with-workload", "l", "WITH_WORKLOAD", FlagDef::BOOL},
    {"component", "c", "COMPONENT", FlagDef::STRING},
    {"cleanup-all", "r", "CLEANUP_ALL", FlagDef::BOOL},
    {"output-dir", "o", "OUTPUT_DIR", FlagDef::STRING},
    {"sleep-interval-seconds", "s", "SLEEP_INTERVAL_SECONDS", FlagDef::INT},
    {"mig-strategy", "m", "MIG_STRATEGY", FlagDef::STRING},
    {"metrics-port", "p", "METRICS_PORT", FlagDef::INT},
    {"default-gpu-workload-config", "g", "DEFAULT_GPU_WORKLOAD_CONFIG", FlagDef::STRING},
    {"disable-dev-char-symlink-creation", "", "DISABLE_DEV_CHAR_SYMLINK_CREATION", FlagDef::BOOL},
};

// strconv.ParseBool
bool parseBool(const std::string &s, bool *out) {
    static const char *t[] = {"1", "t", "T", "TRUE", "true", "True"};
    static const char *f[] = {"0", "f", "F", "FALSE", "false", "False"};
    for (auto x : t)
        if (s == x) return *out = true, true;
    for (auto x : f)
        if (s == x) return *out = false, true;
    return false;
}
bool parseInt(const std::string &s, int *out) {
    if (s.empty()) return false;
    char *end = nullptr;
    errno = 0;
    long v = strtol(s.c_str(), &end, 10);
    if (errno || *end) return false;
    *out = static_cast<int>(v);
    return true;
}

std::string assign(Flags *f, const FlagDef &d, const std::string &value, const char *how) {
    auto bad = [&](const char *type) {
        return sprintfStr("invalid value \"%s\" for %s %s: not a valid %s", value.c_str(), how, d.name, type);
    };
    std::string n = d.name;
    if (d.kind == FlagDef::BOOL) {
        bool b;
        if (!parseBool(value, &b)) return bad("bool");
        if (n == "with-wait") f->withWait = b;
        else if (n == "with-workload") f->withWorkload = b;
        else if (n == "cleanup-all") f->cleanupAll = b;
        else f->disableDevCharSymlinkCreation = b;
    } else if (d.kind == FlagDef::INT) {
        int v;
        if (!parseInt(value, &v)) return bad("int");
        if (n == "sleep-interval-seconds") f->sleepIntervalSeconds = v;
        else f->metricsPort = v;
    } else {
        if (n == "kubeconfig") f->kubeconfig = value;
        else if (n == "node-name") f->nodeName = value;
        else if (n == "namespace") f->namespace_ = value;
        else if (n == "component") f->component = value;
        else if (n == "output-dir") f->outputDir = value;
        else if (n == "mig-strategy") f->migStrategy = value;
        else f->defaultGPUWorkloadConfig = value;
    }
    return "";
}
}  // namespace

std::string parseFlags(int argc, char **argv, Flags *f) {
    bool seen[sizeof(kFlags) / sizeof(kFlags[0])] = {};
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "--") break;
        if (a.size() < 2 || a[0] != '-') return sprintfStr("unexpected argument %s", a.c_str());
        std::string name = a.substr(a[1] == '-' ? 2 : 1), value;
        bool hasValue = false;
        size_t eq = name.find('=');
        if (eq != std::string::npos) {
            value = name.substr(eq + 1);
            name = name.substr(0, eq);
            hasValue = true;
        }
        const FlagDef *d = nullptr;
        size_t idx = 0;
        for (; idx < sizeof(kFlags) / sizeof(kFlags[0]); ++idx)
            if (name == kFlags[idx].name || (*kFlags[idx].alias && name == kFlags[idx].alias)) {
                d = &kFlags[idx];
                break;
            }
        if (!d) return sprintfStr("flag provided but not defined: -%s", name.c_str());
        if (d->kind == FlagDef::BOOL) {
            if (!hasValue) value = "true";
        } else if (!hasValue) {
            if (i + 1 >= argc) return sprintfStr("flag needs an argument: -%s", name.c_str());
            value = argv[++i];
        }
        std::string e = assign(f, *d, value, "flag");
        if (!e.empty()) return e;
        seen[idx] = true;
    }
    for (size_t idx = 0; idx < sizeof(kFlags) / sizeof(kFlags[0]); ++idx) {
        if (seen[idx]) continue;
        const char *v = getenv(kFlags[idx].env);
        if (!v || !*v) continue;
        std::string e = assign(f, kFlags[idx], v, "env");
        if (!e.empty()) return e;
    }
    return "";
}

bool isValidComponent(const std::string &c) {
    static const char *ok[] = {"driver", "toolkit", "cuda", "metrics", "plugin", "mofed",
                               "vfio-pci", "vgpu-manager", "vgpu-devices", "nvidia-fs"};
    for (auto x : ok)
        if (c == x) return true;
    return false;
}

std::string validateFlags(const Flags &f) {
    if (f.component.empty()) return "invalid -c <component-name> flag: must not be empty string";
    if (!isValidComponent(f.component)) return "invalid -c <component-name> flag value: " + f.component;
    if (f.component == "plugin") {
        if (f.nodeName.empty()) return "invalid -n <node-name> flag: must not be empty string for plugin validation";
        if (f.namespace_.empty()) return "invalid -ns <namespace> flag: must not be empty string for plugin validation";
    }
    if (f.component == "cuda" && f.namespace_.empty())
        return "invalid -ns <namespace> flag: must not be empty string for cuda validation";
    if (f.component == "metrics") {
        if (f.metricsPort == defaultMetricsPort)
            return "invalid -p <port> flag: must not be empty or 0 for the metrics component";
        if (f.nodeName.empty()) return "invalid -n <node-name> flag: must not be empty string for metrics exporter";
    }
    if (f.nodeName.empty() && (f.component == "vfio-pci" || f.component == "vgpu-manager" || f.component == "vgpu-devices"))
        return "invalid -n <node-name> flag: must not be empty string for " + f.component + " validation";
    return "";
}

// ------------------------------------------------------------------------------ status files
std::string createStatusFile(const std::string &path) {
    FILE *fp = fopen(path.c_str(), "w");  // os.Create: truncate or create, empty
    if (!fp) return sprintfStr("unable to create status file %s: %s", path.c_str(), strerror(errno));
    fclose(fp);
    return "";
}

std::string deleteStatusFile(const std::string &path) {
    if (unlink(path.c_str()) != 0 && errno != ENOENT)
        return sprintfStr("unable to remove driver status file %s: %s", path.c_str(), strerror(errno));
    return "";  // already removed is fine
}

static int removeAll(const std::string &path) {
    struct stat st;
    if (lstat(path.c_str(), &st) != 0) return errno == ENOENT ? 0 : -1;
    if (S_ISDIR(st.st_mode)) {
        DIR *d = opendir(path.c_str());
        if (!d) return -1;
        while (struct dirent *e = readdir(d)) {
            if (!strcmp(e->d_name, ".") || !strcmp(e->d_name, "..")) continue;
            if (removeAll(path + "/" + e->d_name) != 0) {
                closedir(d);
                return -1;
            }
        }
        closedir(d);
        return rmdir(path.c_str());
    }
    return unlink(path.c_str());
}

// ------------------------------------------------------------------------------ pod spec
namespace {
std::string trim(const std::string &s) {
    size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
    return a == std::string::npos ? "" : s.substr(a, b - a + 1);
}
std::string unquote(std::string s) {
    s = trim(s);
    if (s.size() >= 2 && ((s.front() == '"' && s.back() == '"') || (s.front() == '\'' && s.back() == '\'')))
        s = s.substr(1, s.size() - 2);
    return s;
}
// ['sh', '-c'] or ["vectorAdd"]
std::vector<std::string> parseFlowList(const std::string &v) {
    std::vector<std::string> out;
    std::string s = trim(v);
    if (s.size() < 2 || s.front() != '[' || s.back() != ']') return out;
    s = s.substr(1, s.size() - 2);
    std::string cur;
    char quote = 0;
    for (char ch : s) {
        if (quote) {
            if (ch == quote) quote = 0;
            else cur += ch;
        } else if (ch == '"' || ch == '\'') {
            quote = ch;
        } else if (ch == ',') {
            out.push_back(trim(cur));
            cur.clear();
        } else {
            cur += ch;
        }
    }
    if (!trim(cur).empty() || !out.empty()) out.push_back(trim(cur));
    return out;
}
}  // namespace

// Cuts a trailing comment: '#' opens one only at the start of the line or after white space, and never inside quotes.
static std::string stripComment(const std::string &line) {
    char quote = 0;
    for (size_t i = 0; i < line.size(); ++i) {
        const char ch = line[i];
        if (quote) {
            if (ch == quote) quote = 0;
        } else if (ch == '"' || ch == '\'') {
            quote = ch;
        } else if (ch == '#' && (i == 0 || line[i - 1] == ' ' || line[i - 1] == '\t')) {
            return line.substr(0, i);
        }
    }
    return line;
}

// A line-oriented reader for the block-style subset of YAML that the workload manifests use
// (validator/cuda-workload-validation.yaml, validator/plugin-workload-validation.yaml):
// nested maps by indentation, "- " sequences of maps, flow lists for command/args. What it cannot
// read it refuses by name (see validator.hpp) instead of silently dropping.
std::string loadPodSpec(const std::string &path, PodSpec *pod) {
    std::ifstream in(path);
    if (!in) return sprintfStr("open %s: %s", path.c_str(), strerror(errno));
    std::string line, kind;
    std::string section;      // "initContainers" / "containers" / ""
    std::string sub;          // "env" / "limits" / "labels" / "volumeMounts" / "volumes" ...
    int subIndent = -1;       // indentation of the key that opened `sub`
    std::string pendingEnvName, pendingMountName, pendingVolumeName;
    int sectionIndent = -1;
    std::vector<std::string> *listTarget = nullptr;  // command:/args: given as a block sequence on the next lines
    int listIndent = 0;
    int lineNo = 0;
    bool sawContent = false;
    auto unsupported = [&](const char *what) {
        return sprintfStr("%s:%d: %s is not supported by the validator's manifest reader (it reads plain block-style maps, "
                          "\"- \" sequences and flow lists only)", path.c_str(), lineNo, what);
    };
    while (std::getline(in, line)) {
        ++lineNo;
        line = stripComment(line);
        if (trim(line).empty()) continue;
        if (trim(line) == "---") {
            if (sawContent) return unsupported("a second YAML document (\"---\")");
            continue;
        }
        sawContent = true;
        const size_t firstNonSpace = line.find_first_not_of(' ');
        if (firstNonSpace != std::string::npos && line[firstNonSpace] == '\t') return unsupported("tab indentation");
        int indent = static_cast<int>(firstNonSpace);
        std::string t = trim(line);
        bool dash = t.rfind("- ", 0) == 0;
        if (dash) t = trim(t.substr(2));
        // block-style sequence under command:/args: ("- sh" / "- -c" / "- vectorAdd")
        if (listTarget && dash && indent >= listIndent) {
            listTarget->push_back(unquote(t));
            continue;
        }
        listTarget = nullptr;
        size_t colon = t.find(':');
        if (colon == std::string::npos) continue;
        std::string key = unquote(t.substr(0, colon)), val = trim(t.substr(colon + 1));
        if (key == "<<") return unsupported("a merge key (\"<<\")");
        if (!val.empty()) {
            if (val[0] == '&') return unsupported("an anchor (\"&name\")");
            if (val[0] == '*') return unsupported("an alias (\"*name\")");
            if (val[0] == '{') return unsupported("a flow mapping (\"{...}\")");
            if (val[0] == '|' || val[0] == '>') return unsupported("a block scalar (\"|\" or \">\")");
        }
        // a line at or left of the key that opened the current sub-map closes it (a "- " item may sit at the key's indent)
        if (!sub.empty() && subIndent >= 0 && (indent < subIndent || (indent == subIndent && !dash))) {
            sub.clear();
            subIndent = -1;
        }

        if (indent == 0) {
            section.clear();
            sub = key;
            subIndent = -1;
            if (key == "kind") kind = unquote(val);
            continue;
        }
        if (key == "initContainers" || key == "containers") {
            section = key;
            sectionIndent = indent;
            sub.clear();
            continue;
        }
        if (!section.empty() && indent <= sectionIndent && !dash) section.clear();
        if (section.empty()) {
            if (key == "labels" || key == "tolerations" || key == "volumes") {
                sub = key;
                subIndent = indent;
            } else if (key == "app" && sub == "labels") pod->appLabel = unquote(val);
            else if (sub == "volumes" && key == "name" && dash) pendingVolumeName = unquote(val);
            else if (sub == "volumes" && key == "path" && !pendingVolumeName.empty()) pod->hostPathVolumes[pendingVolumeName] = unquote(val);
            else if (sub == "volumes") continue;  // hostPath: / type: ...
            else if (key == "generateName") pod->generateName = unquote(val);
            else if (key == "namespace") pod->namespace_ = unquote(val);
            else if (key == "restartPolicy") pod->restartPolicy = unquote(val);
            continue;
        }
        if (section != "initContainers") continue;  // the main container only echoes success
        if (key == "env" || key == "resources" || key == "securityContext" || key == "volumeMounts") {
            sub = key;
            subIndent = indent;
            continue;
        }
        if (key == "limits" || key == "requests") {
            sub = key;
            subIndent = indent;
            continue;
        }
        if (sub == "env" && key == "name" && dash) {
            pendingEnvName = unquote(val);
            continue;
        }
        if (sub == "env" && key == "value") {
            if (!pendingEnvName.empty()) pod->initEnv[pendingEnvName] = unquote(val);
            pendingEnvName.clear();
            continue;
        }
        if (sub == "volumeMounts") {
            if (key == "name") pendingMountName = unquote(val);
            else if (key == "mountPath" && !pendingMountName.empty()) pod->initVolumeMounts[pendingMountName] = unquote(val);
            continue;
        }
        if (sub == "limits") {
            pod->initResourceLimits[key] = unquote(val);
            continue;
        }
        if (key == "name" && dash) {
            pod->initName = unquote(val);
            sub.clear();
        } else if (key == "image") pod->image = unquote(val);
        else if (key == "imagePullPolicy") pod->imagePullPolicy = unquote(val);
        else if (key == "command" || key == "args") {
            std::vector<std::string> &dst = key == "command" ? pod->initCommand : pod->initArgs;
            dst = parseFlowList(val);
            if (val.empty()) {  // block sequence follows
                listTarget = &dst;
                listIndent = indent;
            }
        }
    }
    if (kind != "Pod") return sprintfStr("no kind \"Pod\" is registered in %s (kind: %s)", path.c_str(), kind.c_str());
    if (pod->initArgs.empty()) return sprintfStr("%s: no initContainers[0].args", path.c_str());
    return "";
}

bool mountsHostDir(const PodSpec &pod, const std::string &dir) {
    for (auto &m : pod.initVolumeMounts) {
        auto v = pod.hostPathVolumes.find(m.first);
        if (m.second == dir && v != pod.hostPathVolumes.end() && v->second == dir) return true;
    }
    return false;
}

void ensureValidationsVolume(PodSpec *pod) {
    if (mountsHostDir(*pod, defaultStatusPath)) return;
    pod->hostPathVolumes["run-nvidia-validations"] = defaultStatusPath;
    pod->initVolumeMounts["run-nvidia-validations"] = defaultStatusPath;
}

// ------------------------------------------------------------------------------ node resources
namespace {
class LocalNodeResources : public NodeResources {
public:
    std::string get(ResourceList *capacity, ResourceList *allocatable) override {
        capacity->clear();
        allocatable->clear();
        if (const char *file = getenv("GPUPROBE_NODE_CAPACITY_FILE")) {
            std::ifstream in(file);
            if (!in) return sprintfStr("nodes \"%s\" not found (%s: %s)", envOr("NODE_NAME", ""), file, strerror(errno));
            std::string name;
            long long q;
            while (in >> name >> q) (*capacity)[name] = (*allocatable)[name] = q;
            return "";
        }
        int n = 0;
        if (gpuprobe_device_count(&n) != GPUPROBE_OK) n = 0;
        if (n > 0) (*capacity)[genericGPUResourceType] = (*allocatable)[genericGPUResourceType] = n;
        return "";
    }
};

std::string randomSuffix() {
    static const char alnum[] = "bcdfghjklmnpqrstvwxz2456789";  // the alphabet k8s generateName uses
    std::random_device rd;
    std::string s;
    for (int i = 0; i < 5; ++i) s += alnum[rd() % (sizeof(alnum) - 1)];
    return s;
}

class LocalLauncher : public WorkloadLauncher {
public:
    explicit LocalLauncher(const Flags &f) : flags_(f) {}
    ~LocalLauncher() override { killChild(); }

    std::string create(const PodSpec &pod, std::string *name) override {
        killChild();  // "check if ... validation pod is already running and cleanup" (main.go:1262-1279)
        pod_ = pod;
        name_ = pod.generateName + randomSuffix();
        *name = name_;
        restarts_ = 0;
        succeeded_ = false;
        return spawn();
    }

    std::string phase(const std::string &name, std::string *phase) override {
        if (name != name_) return sprintfStr("pods \"%s\" not found", name.c_str());
        if (succeeded_) return *phase = "Succeeded", "";
        if (pid_ > 0) {
            int status = 0;
            pid_t r = waitpid(pid_, &status, WNOHANG);
            if (r == 0) return *phase = "Pending", "";  // init container still running
            pid_ = -1;
            if (r > 0 && WIFEXITED(status) && WEXITSTATUS(status) == 0) {
                succeeded_ = true;  // init container done; the main container only echoes success
                return *phase = "Succeeded", "";
            }
            lastExit_ = (r > 0 && WIFEXITED(status)) ? WEXITSTATUS(status) : -1;
            if (pod_.restartPolicy != "OnFailure" && pod_.restartPolicy != "Always") return *phase = "Failed", "";
            ++restarts_;
            retryAt_ = std::chrono::steady_clock::now() + std::chrono::seconds(std::max(0, flags_.sleepIntervalSeconds));
            logInfo("init container %s of pod %s exited with code %d; restart %d (restartPolicy: %s)", pod_.initName.c_str(),
                    name_.c_str(), lastExit_, restarts_, pod_.restartPolicy.c_str());
        }
        if (std::chrono::steady_clock::now() >= retryAt_) {
            std::string e = spawn();
            if (!e.empty()) return e;
        }
        return *phase = "Pending", "";
    }

private:
    std::string spawn() {
        std::vector<std::string> argvS = pod_.initCommand;
        if (argvS.empty()) argvS = {"sh", "-c"};
        for (auto &a : pod_.initArgs) argvS.push_back(a);
        std::vector<std::string> envS;
        for (char **e = environ; *e; ++e) envS.push_back(*e);
        auto setEnv = [&](const std::string &k, const std::string &v) {
            for (auto &e : envS)
                if (e.rfind(k + "=", 0) == 0) {
                    e = k + "=" + v;
                    return;
                }
            envS.push_back(k + "=" + v);
        };
        for (auto &kv : pod_.initEnv) setEnv(kv.first, kv.second);
        // The pod has its own filesystem: what it writes under /run/nvidia/validations reaches this validator's output
        // directory only through a hostPath mount of that path. Without one the child gets a directory of its own.
        setEnv("OUTPUT_DIR", mountsHostDir(pod_, defaultStatusPath) ? flags_.outputDir : "/nonexistent/pod-private" + flags_.outputDir);
        if (!flags_.nodeName.empty()) setEnv("NODE_NAME", flags_.nodeName);
        // a pod that was allocated one GPU by the device plugin sees one GPU
        for (auto &kv : pod_.initResourceLimits)
            if (kv.first.rfind("nvidia.com/", 0) == 0 && !getenv("CUDA_VISIBLE_DEVICES")) setEnv("CUDA_VISIBLE_DEVICES", "0");
        if (const char *bin = getenv("GPUPROBE_BIN_DIR")) setEnv("PATH", std::string(bin) + ":" + envOr("PATH", "/usr/bin:/bin"));

        pid_t pid = fork();
        if (pid < 0) return sprintfStr("failed to create %s pod %s, err fork: %s", pod_.appLabel.c_str(), name_.c_str(), strerror(errno));
        if (pid == 0) {
            std::vector<char *> av, ev;
            for (auto &s : argvS) av.push_back(const_cast<char *>(s.c_str()));
            av.push_back(nullptr);
            for (auto &s : envS) ev.push_back(const_cast<char *>(s.c_str()));
            ev.push_back(nullptr);
            execvpe(av[0], av.data(), ev.data());
            _exit(127);
        }
        pid_ = pid;
        return "";
    }
    void killChild() {
        if (pid_ > 0) {
            kill(pid_, SIGKILL);
            waitpid(pid_, nullptr, 0);
            pid_ = -1;
        }
    }
    Flags flags_;
    PodSpec pod_;
    std::string name_;
    pid_t pid_ = -1;
    int restarts_ = 0, lastExit_ = 0;
    bool succeeded_ = false;
    std::chrono::steady_clock::time_point retryAt_{};
};
}  // namespace

std::unique_ptr<NodeResources> makeLocalNodeResources() { return std::make_unique<LocalNodeResources>(); }
std::unique_ptr<WorkloadLauncher> makeLocalLauncher(const Flags &f) { return std::make_unique<LocalLauncher>(f); }

std::string waitForPod(WorkloadLauncher &l, const std::string &name, int retries, int pollMillis) {
    for (int i = 0; i < retries; ++i) {
        std::string phase;
        std::string e = l.phase(name, &phase);
        if (!e.empty()) return sprintfStr("failed to get pod %s, err %s", name.c_str(), e.c_str());
        if (phase != "Succeeded") {
            logInfo("pod %s is curently in %s phase", name.c_str(), phase.c_str());  // sic, main.go:1051
            std::this_thread::sleep_for(std::chrono::milliseconds(pollMillis));
            continue;
        }
        logInfo("pod %s have run successfully", name.c_str());
        return "";
    }
    return sprintfStr("gave up waiting for pod %s to be available", name.c_str());
}

// ------------------------------------------------------------------------------ plugin resources
static bool hasPrefix(const std::string &s, const char *p) { return s.rfind(p, 0) == 0; }

std::string availableMIGResourceName(const ResourceList &r) {
    for (auto &kv : r)
        if (hasPrefix(kv.first, migGPUResourcePrefix) && kv.second >= 1) return kv.first;
    return "";
}
std::string availableGenericResourceName(const ResourceList &r) {
    for (auto &kv : r)
        if (hasPrefix(kv.first, genericGPUResourceType) && kv.second >= 1) return kv.first;
    return "";
}
int64_t countGPUResources(const ResourceList &capacity) {
    int64_t count = 0;
    for (auto &kv : capacity)
        if (hasPrefix(kv.first, migGPUResourcePrefix) || hasPrefix(kv.first, genericGPUResourceType)) count += kv.second;
    return count;
}
std::string getGPUResourceName(const ResourceList &allocatable, std::string *name) {
    // use mig resource if one is available to run workload (main.go:1151-1158)
    *name = availableMIGResourceName(allocatable);
    if (!name->empty()) return "";
    *name = availableGenericResourceName(allocatable);
    if (!name->empty()) return "";
    return "Unable to find any allocatable GPU resource";
}

// ------------------------------------------------------------------------------ components
static std::string manifestPath(const Runtime &rt, const char *referencePath) {
    if (rt.manifestDir.empty()) return referencePath;
    const char *base = strrchr(referencePath, '/');
    return rt.manifestDir + "/" + (base ? base + 1 : referencePath);
}

// Shared body of CUDA.runWorkload (main.go:1217-1293) and Plugin.runWorkload (:931-1015).
static std::string runWorkload(Runtime &rt, const char *specPath, const char *what, const std::string &resourceName) {
    PodSpec pod;
    std::string e = loadPodSpec(manifestPath(rt, specPath), &pod);
    if (!e.empty()) return e;
    pod.namespace_ = rt.flags.namespace_;
    pod.image = envOr("VALIDATOR_IMAGE", "");
    if (*envOr("VALIDATOR_IMAGE_PULL_POLICY", "")) pod.imagePullPolicy = envOr("VALIDATOR_IMAGE_PULL_POLICY", "");
    if (*envOr("VALIDATOR_IMAGE_PULL_SECRETS", "")) {
        std::stringstream ss(envOr("VALIDATOR_IMAGE_PULL_SECRETS", ""));
        std::string s;
        while (std::getline(ss, s, ',')) pod.imagePullSecrets.push_back(s);
    }
    if (*envOr("VALIDATOR_RUNTIME_CLASS", "")) pod.runtimeClassName = envOr("VALIDATOR_RUNTIME_CLASS", "");
    // setOwnerReference / setTolerations (main.go:1017-1039) copy metadata from the validator DaemonSet: API-only.
    pod.nodeName = rt.flags.nodeName;
    if (!resourceName.empty()) pod.initResourceLimits = {{resourceName, "1"}};
    // New (SURVEY.md section 5 "Subtlety"): validator.cuda.env lands on the DaemonSet's init container, not on
    // the workload pod, so the probe knobs are forwarded explicitly next to the image/pull settings.
    for (char **ev = environ; *ev; ++ev)
        if (!strncmp(*ev, "GPUPROBE_", 9)) {
            const char *eq = strchr(*ev, '=');
            if (eq) pod.initEnv[std::string(*ev, eq - *ev)] = eq + 1;
        }
    // the plugin workload (one allocated GPU) must not overwrite the cuda workload's node-wide results
    if (!resourceName.empty()) pod.initEnv["GPUPROBE_RESULTS_BASENAME"] = "gpuprobe-plugin-results";
    // New: the results file has to leave the pod. The reference's workload manifests declare no volumes, so the pod gets the
    // validations hostPath here if its manifest does not already mount it (GPUPROBE_NO_VOLUME_INJECT=1: tests of the
    // failure mode only).
    if (!*envOr("GPUPROBE_NO_VOLUME_INJECT", "")) ensureValidationsVolume(&pod);
    if (!mountsHostDir(pod, defaultStatusPath))
        logInfo("warning: %s workload pod does not mount %s: its probe results will not reach the metrics exporter", what,
                defaultStatusPath);
    std::string name;
    e = rt.launcher->create(pod, &name);
    if (!e.empty()) return sprintfStr("failed to create %s validation pod %s, err %s", what, pod.generateName.c_str(), e.c_str());
    int retries = podCreationWaitRetries;
    if (!rt.keepRetryCounts && rt.pollMillis > 0 && rt.pollMillis < podCreationSleepIntervalSeconds * 1000)
        retries = podCreationWaitRetries * (podCreationSleepIntervalSeconds * 1000 / rt.pollMillis);
    return waitForPod(*rt.launcher, name, retries, rt.pollMillis);
}

std::string validateCUDA(Runtime &rt) {
    std::string e = deleteStatusFile(rt.flags.outputDir + "/" + cudaStatusFile);
    if (!e.empty()) return e;
    e = runWorkload(rt, cudaWorkloadPodSpecPath, "cuda", "");
    if (!e.empty()) return e;
    return createStatusFile(rt.flags.outputDir + "/" + cudaStatusFile);
}

static std::string validateGPUResource(Runtime &rt) {
    int retries = gpuResourceDiscoveryWaitRetries;
    if (!rt.keepRetryCounts && rt.discoveryPollMillis > 0 && rt.discoveryPollMillis < gpuResourceDiscoveryIntervalSeconds * 1000)
        retries = gpuResourceDiscoveryWaitRetries * (gpuResourceDiscoveryIntervalSeconds * 1000 / rt.discoveryPollMillis);
    for (int retry = 1; retry <= retries; ++retry) {
        ResourceList cap, alloc;
        std::string e = rt.node->get(&cap, &alloc);
        if (!e.empty())
            return sprintfStr("unable to fetch node by name %s to check for GPU resources: %s", rt.flags.nodeName.c_str(), e.c_str());
        if (!availableMIGResourceName(cap).empty()) return "";
        if (!availableGenericResourceName(cap).empty()) return "";
        logInfo("GPU resources are not yet discovered by the node, retry: %d", retry);
        std::this_thread::sleep_for(std::chrono::milliseconds(rt.discoveryPollMillis));
    }
    return "GPU resources are not discovered by the node";
}

std::string validatePlugin(Runtime &rt) {
    std::string e = deleteStatusFile(rt.flags.outputDir + "/" + pluginStatusFile);
    if (!e.empty()) return e;
    e = validateGPUResource(rt);
    if (!e.empty()) return e;
    if (rt.flags.withWorkload) {
        ResourceList cap, alloc;
        e = rt.node->get(&cap, &alloc);
        if (!e.empty())
            return sprintfStr("unable to fetch node by name %s to check for GPU resources: %s", rt.flags.nodeName.c_str(), e.c_str());
        std::string resourceName;
        e = getGPUResourceName(alloc, &resourceName);
        if (!e.empty()) return e;
        e = runWorkload(rt, pluginWorkloadPodSpecPath, "plugin", resourceName);
        if (!e.empty()) return e;
    }
    return createStatusFile(rt.flags.outputDir + "/" + pluginStatusFile);
}

// runCommand / runCommandWithWait (main.go:547-572)
static int runCommand(const std::vector<std::string> &argvS, bool silent) {
    pid_t pid = fork();
    if (pid < 0) return -1;
    if (pid == 0) {
        if (silent) {
            FILE *n1 = freopen("/dev/null", "w", stdout);
            FILE *n2 = freopen("/dev/null", "w", stderr);
            (void)n1;
            (void)n2;
        }
        std::vector<char *> av;
        for (auto &s : argvS) av.push_back(const_cast<char *>(s.c_str()));
        av.push_back(nullptr);
        execvp(av[0], av.data());
        _exit(127);
    }
    int status = 0;
    waitpid(pid, &status, 0);
    return WIFEXITED(status) ? WEXITSTATUS(status) : -1;
}
static std::string runCommandMaybeWait(const Runtime &rt, const std::vector<std::string> &argvS, bool silent) {
    for (;;) {
        if (rt.flags.withWait) {
            std::string shown;
            for (size_t i = 1; i < argvS.size(); ++i) shown += (i > 1 ? " " : "") + argvS[i];
            printf("running command %s with args [%s]\n", argvS[0].c_str(), shown.c_str());
        }
        int rc = runCommand(argvS, silent);
        if (rc == 0) return "";
        if (!rt.flags.withWait) return sprintfStr("exit status %d", rc);
        printf("command failed, retrying after %d seconds\n", rt.flags.sleepIntervalSeconds);
        std::this_thread::sleep_for(std::chrono::seconds(rt.flags.sleepIntervalSeconds));
    }
}

// Driver.runValidation (main.go:609-628): getDriverRoot (:585-593) -- a pre-installed host driver is seen at /host, else
// the driver container's root /run/nvidia/driver; for a driver container first assertDriverContainerReady (:598-607:
// `stat /run/nvidia/validations/.driver-ctr-ready`), then `chroot <root> nvidia-smi`.
// GPUPROBE_HEADLESS=1 (tools/full_pass.py, the GPU tests) adds what the reference does not have: on a box with
// neither mount the driver in the current root is checked with a bare `nvidia-smi` and counts as a host driver.
// Without that variable the mirror does what the reference would do there -- and fails the same way.
static std::string runDriverValidation(const Runtime &rt, bool silent, bool *isHostDriver) {
    std::string driverRoot = "/run/nvidia/driver";
    *isHostDriver = false;
    struct stat st;
    if (lstat("/host/usr/bin/nvidia-smi", &st) == 0 && st.st_size != 0) {
        logInfo("Detected pre-installed driver on the host");
        driverRoot = "/host";
        *isHostDriver = true;
    } else if (*envOr("GPUPROBE_HEADLESS", "") == '1' && !fileExists("/run/nvidia/driver/usr/bin/nvidia-smi")) {
        *isHostDriver = true;
        return runCommandMaybeWait(rt, {"nvidia-smi"}, silent);
    }
    if (!*isHostDriver) {
        logInfo("Driver is not pre-installed on the host. Checking driver container status.");
        std::string e = runCommandMaybeWait(rt, {"bash", "-c", std::string("stat ") + defaultStatusPath + "/.driver-ctr-ready"}, silent);
        if (!e.empty()) return "error checking driver container status: " + e;
    }
    return runCommandMaybeWait(rt, {"chroot", driverRoot, "nvidia-smi"}, silent);
}

std::string validateDriver(Runtime &rt) {
    std::string e = deleteStatusFile(rt.flags.outputDir + "/" + driverStatusFile);
    if (!e.empty()) return e;
    e = deleteStatusFile(rt.flags.outputDir + "/" + hostDriverStatusFile);
    if (!e.empty()) return e;
    bool isHostDriver = false;
    e = runDriverValidation(rt, false, &isHostDriver);
    if (!e.empty()) {
        logError("driver is not ready");
        return e;
    }
    // createDevCharSymlinks (main.go:693-717) needs the privileged /host-dev-char mount: not part of this path.
    return createStatusFile(rt.flags.outputDir + "/" + (isHostDriver ? hostDriverStatusFile : driverStatusFile));
}

std::string validateToolkit(Runtime &rt) {
    std::string e = deleteStatusFile(rt.flags.outputDir + "/" + toolkitStatusFile);
    if (!e.empty()) return e;
    e = runCommandMaybeWait(rt, {"nvidia-smi"}, false);
    if (!e.empty()) {
        printf("toolkit is not ready\n");
        return e;
    }
    return createStatusFile(rt.flags.outputDir + "/" + toolkitStatusFile);
}

// ------------------------------------------------------------------------------ metrics
static void gauge(std::ostringstream &o, const char *name, const char *help, const std::string &node, double v) {
    o << "# HELP " << name << " " << help << "\n# TYPE " << name << " gauge\n"
      << name << "{node=\"" << node << "\"} " << sprintfStr("%.10g", v) << "\n";
}

// NodeMetrics (metrics.go:52-70): the gauge values. Watcher threads write them on the reference's cadence; a scrape
// only formats them, so nothing slow (nvidia-smi, lspci, a file read) ever sits on the scrape path.
namespace {
struct NodeMetricsState {
    std::mutex mu;
    double metricsReady = 0;
    double driverReady = 0, toolkitReady = 0, pluginReady = 0, cudaReady = 0;
    double driverValidation = 0, driverValidationLastSuccess = 0;
    double deviceCount = 0, pluginValidationLastSuccess = 0;
    double nvidiaPciDevices = 0;
    std::string probeProm;  // gpuprobe-results.prom with @NODE@ filled in
};
NodeMetricsState g_nm;

// statusFileCheckDelaySeconds = 30, driverValidationCheckDelaySeconds = 60, pluginValidationCheckDelaySeconds = 30
// (metrics.go:40-47). GPUPROBE_WATCH_MS_PER_SECOND (tests) scales a "second" of those delays, default 1000 ms.
void watchSleep(int seconds) {
    const int msPerSecond = std::max(1, atoi(envOr("GPUPROBE_WATCH_MS_PER_SECOND", "1000")));
    std::this_thread::sleep_for(std::chrono::milliseconds(static_cast<long long>(seconds) * msPerSecond));
}

// watchStatusFile (metrics.go:159-190): stat every 30 s, flip the gauge and log on a transition.
void watchStatusFile(const Flags f, double NodeMetricsState::*gaugeField, const char *statusFile) {
    logInfo("metrics: StatusFile: watching %s", statusFile);
    bool prevReady = false;
    for (;;) {
        bool ready = fileExists(f.outputDir + "/" + statusFile);
        if (!ready && !strcmp(statusFile, driverStatusFile)) ready = fileExists(f.outputDir + "/" + hostDriverStatusFile);  // :169-173
        if (ready != prevReady) {
            prevReady = ready;
            logInfo("metrics: StatusFile: '%s' is %s", statusFile, ready ? "ready" : "not ready");
            std::lock_guard<std::mutex> lock(g_nm.mu);
            g_nm.*gaugeField = ready ? 1.0 : 0.0;
        }
        watchSleep(30);
    }
}

// watchDriverValidation (metrics.go:237-250): Driver.runValidation(silent) every 60 s.
void watchDriverValidation(const Runtime *rt) {
    for (;;) {
        bool isHostDriver = false;
        Runtime once;  // no wait loop inside the watcher: one attempt per period
        once.flags = rt->flags;
        once.flags.withWait = false;
        const bool ok = runDriverValidation(once, true, &isHostDriver).empty();
        {
            std::lock_guard<std::mutex> lock(g_nm.mu);
            g_nm.driverValidation = ok ? 1.0 : 0.0;
            if (ok) g_nm.driverValidationLastSuccess = static_cast<double>(time(nullptr));
        }
        watchSleep(60);
    }
}

// watchDevicePluginValidation (metrics.go:192-235): countGPUResources every 30 s, log when the count changes.
void watchDevicePluginValidation(const Flags f) {
    logInfo("metrics: DevicePlugin validation: node name is %s", f.nodeName.c_str());
    auto node = makeLocalNodeResources();
    long long prevCount = -2;
    for (;;) {
        ResourceList cap, alloc;
        const std::string e = node->get(&cap, &alloc);
        long long count = -1;
        {
            std::lock_guard<std::mutex> lock(g_nm.mu);
            if (!e.empty()) {
                g_nm.deviceCount = -1;
            } else {
                count = countGPUResources(cap);
                g_nm.deviceCount = static_cast<double>(count);
                if (count != 0) g_nm.pluginValidationLastSuccess = static_cast<double>(time(nullptr));
            }
        }
        if (prevCount != count) {
            if (!e.empty()) logError("metrics: DevicePlugin validation: could not list the DevicePlugin devices: %s", e.c_str());
            else logInfo("metrics: DevicePlugin validation: found %lld GPUs exposed by the DevicePlugin", count);
        }
        prevCount = count;
        watchSleep(30);
    }
}

// watchNVIDIAPCI (metrics.go:252-301): `lspci`, count the lines that mention nvidia, every 60 s.
void watchNVIDIAPCI() {
    int prev = -2;
    for (;;) {
        int devCount = -1;
        if (FILE *p = popen("lspci 2>/dev/null", "r")) {
            char line[1024];
            int n = 0;
            while (fgets(line, sizeof(line), p)) {
                for (char *c = line; *c; ++c) *c = static_cast<char>(tolower(static_cast<unsigned char>(*c)));
                if (strstr(line, "nvidia")) ++n;
            }
            if (pclose(p) == 0) devCount = n;
        }
        if (prev != devCount) {
            if (devCount < 0) logError("metrics: PCI devices: Error running 'lspci'");
            else logInfo("metrics: PCI devices: found %d NVIDIA device%s", devCount, devCount > 1 ? "s" : "");
        }
        prev = devCount;
        {
            std::lock_guard<std::mutex> lock(g_nm.mu);
            g_nm.nvidiaPciDevices = devCount;
        }
        watchSleep(60);
    }
}

// New (integration/validator/gpuprobe_metrics.go watchProbeResults): the probe's figures, re-read every 30 s.
void watchProbeResults(const Flags f) {
    const std::string path = f.outputDir + "/" + probeResultsProm;
    logInfo("metrics: gpuprobe: watching %s", path.c_str());
    for (;;) {
        std::ifstream prom(path);
        std::string text, line;
        while (prom && std::getline(prom, line)) {
            size_t at;
            while ((at = line.find("@NODE@")) != std::string::npos) line.replace(at, 6, f.nodeName);
            text += line + "\n";
        }
        {
            std::lock_guard<std::mutex> lock(g_nm.mu);
            g_nm.probeProm = text;
        }
        watchSleep(30);
    }
}
}  // namespace

std::string renderMetrics(const Flags &f, double metricsReadyTs) {
    std::ostringstream o;
    const std::string &node = f.nodeName;
    std::lock_guard<std::mutex> lock(g_nm.mu);
    // names and help strings: validator/metrics.go:77-155
    gauge(o, "gpu_operator_node_metrics_ready_ts_seconds",
          "timestamp (in seconds) of the launch time of the GPU Operator node metrics Pod", node, metricsReadyTs);
    gauge(o, "gpu_operator_node_driver_ready",
          "1 if the driver synchronization barrier of the the local node is open, 0 otherwise", node, g_nm.driverReady);
    gauge(o, "gpu_operator_node_toolkit_ready",
          "1 if the container-toolkit synchronization barrier on the local node is open, 0 otherwise", node, g_nm.toolkitReady);
    gauge(o, "gpu_operator_node_plugin_ready",
          "1 if the device plugin synchronization barrier on the local is open, 0 otherwise", node, g_nm.pluginReady);
    gauge(o, "gpu_operator_node_cuda_ready", "1 if cuda synchronization barrier on the local node is open, 0 otherwise", node,
          g_nm.cudaReady);
    gauge(o, "gpu_operator_node_driver_validation", "1 if the driver validation test passed on the local node, 0 otherwise", node,
          g_nm.driverValidation);
    gauge(o, "gpu_operator_node_driver_validation_last_success_ts_seconds",
          "timestamp (in seconds) of the last successful driver test validation", node, g_nm.driverValidationLastSuccess);
    gauge(o, "gpu_operator_node_device_plugin_devices_total",
          "number of GPU devices exposed by the DevicePlugin on the local node. -1 if failing to inspect the node spec.", node,
          g_nm.deviceCount);
    gauge(o, "gpu_operator_node_device_plugin_validation_last_success_ts_seconds",
          "timestamp (in seconds) of the last time GPU devices were found on the local node spec", node, g_nm.pluginValidationLastSuccess);
    gauge(o, "gpu_operator_nvidia_pci_devices_total", "number of NVIDIA devices found in the node. -1 if failing to count", node,
          g_nm.nvidiaPciDevices);
    // new gauges: the probe's achieved bandwidth, written by the workload next to the ready files
    o << g_nm.probeProm;
    return o.str();
}

std::string runMetrics(Runtime &rt) {
    const double readyTs = static_cast<double>(time(nullptr));
    int srv = socket(AF_INET, SOCK_STREAM, 0);
    if (srv < 0) return sprintfStr("socket: %s", strerror(errno));
    int one = 1;
    setsockopt(srv, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in addr{};
    addr.sin_family = AF_INET;
    addr.sin_addr.s_addr = htonl(INADDR_ANY);
    addr.sin_port = htons(static_cast<uint16_t>(rt.flags.metricsPort));
    if (bind(srv, reinterpret_cast<sockaddr *>(&addr), sizeof(addr)) != 0 || listen(srv, 16) != 0) {
        std::string e = sprintfStr("listen tcp :%d: %s", rt.flags.metricsPort, strerror(errno));
        close(srv);
        return e;
    }
    // NodeMetrics.Run (metrics.go:304-320): stamp the start time, start the watchers, serve.
    std::thread(watchStatusFile, rt.flags, &NodeMetricsState::driverReady, driverStatusFile).detach();
    std::thread(watchStatusFile, rt.flags, &NodeMetricsState::toolkitReady, toolkitStatusFile).detach();
    std::thread(watchStatusFile, rt.flags, &NodeMetricsState::pluginReady, pluginStatusFile).detach();
    std::thread(watchStatusFile, rt.flags, &NodeMetricsState::cudaReady, cudaStatusFile).detach();
    std::thread(watchDriverValidation, &rt).detach();
    std::thread(watchDevicePluginValidation, rt.flags).detach();
    std::thread(watchNVIDIAPCI).detach();
    std::thread(watchProbeResults, rt.flags).detach();
    logInfo("Running the metrics server, listening on :%d/metrics", rt.flags.metricsPort);
    for (;;) {
        int c = accept(srv, nullptr, nullptr);
        if (c < 0) {
            if (errno == EINTR) continue;
            std::string e = sprintfStr("accept: %s", strerror(errno));
            close(srv);
            return e;
        }
        char req[2048];
        ssize_t n = recv(c, req, sizeof(req) - 1, 0);
        req[n > 0 ? n : 0] = 0;
        std::string body, status = "200 OK";
        if (!strncmp(req, "GET /metrics", 12)) {
            body = renderMetrics(rt.flags, readyTs);
        } else {
            status = "404 Not Found";
            body = "404 page not found\n";
        }
        std::string resp = "HTTP/1.1 " + status + "\r\nContent-Type: text/plain; version=0.0.4; charset=utf-8\r\nContent-Length: " +
                           std::to_string(body.size()) + "\r\nConnection: close\r\n\r\n" + body;
        size_t off = 0;
        while (off < resp.size()) {
            ssize_t w = send(c, resp.data() + off, resp.size() - off, MSG_NOSIGNAL);
            if (w <= 0) break;
            off += static_cast<size_t>(w);
        }
        close(c);
    }
}

// ------------------------------------------------------------------------------ start
std::string start(Runtime &rt) {
    const Flags &f = rt.flags;
    if (f.cleanupAll) {
        if (removeAll(f.outputDir) != 0 && errno != ENOENT) return sprintfStr("unlinkat %s: %s", f.outputDir.c_str(), strerror(errno));
    }
    if (mkdir(f.outputDir.c_str(), 0755) != 0 && errno != EEXIST)
        return sprintfStr("mkdir %s: %s", f.outputDir.c_str(), strerror(errno));

    std::string e;
    if (f.component == "driver") {
        e = validateDriver(rt);
        return e.empty() ? "" : "error validating driver installation: " + e;
    }
    if (f.component == "toolkit") {
        e = validateToolkit(rt);
        return e.empty() ? "" : "error validating toolkit installation: " + e;
    }
    if (f.component == "cuda") {
        e = validateCUDA(rt);
        return e.empty() ? "" : "error validating cuda workload: " + e;
    }
    if (f.component == "plugin") {
        e = validatePlugin(rt);
        return e.empty() ? "" : "error validating plugin installation: " + e;
    }
    if (f.component == "metrics") {
        e = runMetrics(rt);
        return e.empty() ? "" : "error running validation-metrics exporter: " + e;
    }
    // mofed, nvidia-fs, vfio-pci, vgpu-manager, vgpu-devices: valid component names (main.go:372-397) that
    // lie outside the CUDA-workload path this repo rebuilds.
    return "component " + f.component + " is not part of the CUDA-workload path and is not built here";
}

}  // namespace validator
