// kungfu-run: per-host launcher (static and elastic watch mode).
// Parity: /root/reference/srcs/go/cmd/kungfu-run/app/kungfu-run.go:18-112.
#include "elastic/runner.hpp"

int main(int argc, char **argv) { return kf::kungfu_run_main(argc, argv); }
