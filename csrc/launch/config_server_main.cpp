// kungfu-config-server: standalone REST store of the cluster config.
// Parity: /root/reference/srcs/go/cmd/kungfu-config-server/kungfu-config-server.go:19-65.
#include "elastic/runner.hpp"

int main(int argc, char **argv) { return kf::config_server_main(argc, argv); }
