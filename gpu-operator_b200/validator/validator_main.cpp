// validator_main.cpp -- `nvidia-validator`: same command line, env, status files and exit
// behaviour as the reference binary (validator/main.go:207-337) for the components on the
// CUDA-workload path. Extra (headless) knobs, all optional:
//   GPUPROBE_MANIFEST_DIR   directory holding cuda-/plugin-workload-validation.yaml
//                           (default: the reference's /var/nvidia/manifests)
//   GPUPROBE_POLL_MS        pod-phase poll period in ms (reference: 5000, main.go:161)
//   GPUPROBE_KEEP_RETRY_COUNTS=1  with a shorter poll, keep 60/30 retries instead of the 300 s/150 s budgets
//   GPUPROBE_BIN_DIR        prepended to the workload's PATH so that `vectorAdd` resolves
#include <signal.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "validator.hpp"

using namespace validator;

static void onSignal(int sig) {
    // handleSignal (main.go:329-337): log.Fatalf -> exit status 1
    const char *name = sig == SIGTERM ? "terminated" : sig == SIGHUP ? "hangup" : sig == SIGINT ? "interrupt" : "quit";
    char buf[160];
    int n = snprintf(buf, sizeof(buf), "Exiting due to signal [%s] notification for pid [%d]\n", name, getpid());
    if (n > 0) (void)!write(2, buf, static_cast<size_t>(n));
    _exit(1);
}

int main(int argc, char **argv) {
    for (int s : {SIGTERM, SIGHUP, SIGINT, SIGQUIT}) signal(s, onSignal);
    Runtime rt;
    std::string err = parseFlags(argc, argv, &rt.flags);
    if (err.empty()) err = validateFlags(rt.flags);  // cli.App.Before
    if (err.empty()) {
        rt.launcher = makeLocalLauncher(rt.flags);
        rt.node = makeLocalNodeResources();
        if (const char *d = getenv("GPUPROBE_MANIFEST_DIR")) rt.manifestDir = d;
        if (const char *p = getenv("GPUPROBE_POLL_MS")) {
            int ms = atoi(p);
            if (ms > 0) rt.pollMillis = rt.discoveryPollMillis = ms;
        }
        if (const char *k = getenv("GPUPROBE_KEEP_RETRY_COUNTS")) rt.keepRetryCounts = (*k == '1');
        err = start(rt);
    }
    if (!err.empty()) {
        fprintf(stderr, "Error: %s\n", err.c_str());  // main.go:321-326
        return 1;
    }
    return 0;
}
