// kungfu-rrun / kungfu-distribute: SSH fan-out helpers.
//
//   kungfu-distribute -H ip:slots,... [-u user] <command...>     run a command on every host
//   kungfu-rrun -H ip:slots,... -np N [-u user] [kungfu-run flags] <prog> [args]
//                                                   start `kungfu-run` on every host of the job
// Parity: /root/reference/srcs/go/cmd/kungfu-rrun/rrun.go,
// /root/reference/srcs/go/cmd/kungfu-distribute/kungfu-distribute.go:76-88,
// /root/reference/srcs/go/utils/runner/remote/remote.go:22-131, /root/reference/srcs/go/utils/ssh/ssh.go.
// The reference embeds an SSH client; here the system `ssh` binary is driven
// (BatchMode, no tty), one child per host, output prefixed per host.
#include <sys/wait.h>
#include <unistd.h>

#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "core/plan.hpp"

using namespace kf;

static std::string shell_quote(const std::string &s)
{
    std::string o = "'";
    for (char c : s) {
        if (c == '\'') o += "'\\''";
        else o += c;
    }
    return o + "'";
}

static int run_on(const std::string &user, const HostSpec &h, const std::string &cmd, bool dry)
{
    const std::string target = (user.empty() ? "" : user + "@") + h.public_addr;
    if (dry) {
        printf("ssh -o BatchMode=yes %s %s\n", target.c_str(), shell_quote(cmd).c_str());
        return 0;
    }
    int pfd[2];
    if (pipe(pfd) != 0) return 1;
    pid_t pid = fork();
    if (pid == 0) {
        dup2(pfd[1], 1);
        dup2(pfd[1], 2);
        close(pfd[0]);
        execlp("ssh", "ssh", "-o", "BatchMode=yes", "-o", "StrictHostKeyChecking=no", target.c_str(), cmd.c_str(), (char *)nullptr);
        fprintf(stderr, "cannot exec ssh: %s\n", strerror(errno));
        _exit(127);
    }
    close(pfd[1]);
    FILE *in   = fdopen(pfd[0], "r");
    char *line = nullptr;
    size_t cap = 0;
    while (getline(&line, &cap, in) > 0) printf("[%s] %s", h.public_addr.c_str(), line);
    free(line);
    fclose(in);
    int st = 0;
    waitpid(pid, &st, 0);
    return WIFEXITED(st) ? WEXITSTATUS(st) : 1;
}

int main(int argc, char **argv)
{
    const bool rrun = strstr(argv[0], "rrun") != nullptr;
    std::string hosts = "127.0.0.1:1", user, self_flag;
    bool dry = false;
    std::vector<std::string> pass;  // flags forwarded to kungfu-run
    int i = 1;
    for (; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "-H" && i + 1 < argc) hosts = argv[++i];
        else if (a == "-u" && i + 1 < argc) user = argv[++i];
        else if (a == "-dry-run") dry = true;
        else if (rrun && a.size() > 1 && a[0] == '-') {
            pass.push_back(a);
            if (a.find('=') == std::string::npos && a != "-w" && a != "-k" && a != "-q" && a != "-allow-nvlink" && i + 1 < argc)
                pass.push_back(argv[++i]);
        } else break;
    }
    if (i >= argc) {
        fprintf(stderr, "usage: %s -H ip:slots[:public],... [-u user] [-dry-run] %s\n", argv[0],
                rrun ? "[kungfu-run flags] <prog> [args]" : "<command...>");
        return 2;
    }
    HostList hl;
    if (!parse_host_list(hosts, &hl)) {
        fprintf(stderr, "bad -H %s\n", hosts.c_str());
        return 2;
    }
    std::string tail;
    for (; i < argc; ++i) tail += std::string(tail.empty() ? "" : " ") + shell_quote(argv[i]);
    std::vector<int> rc(hl.size(), 0);
    std::vector<std::thread> ths;
    for (size_t k = 0; k < hl.size(); ++k) {
        std::string cmd = tail;
        if (rrun) {
            cmd = "kungfu-run -H " + shell_quote(hosts) + " -self " + format_ipv4(hl[k].ipv4);
            for (auto &f : pass) cmd += " " + shell_quote(f);
            cmd += " " + tail;
        }
        ths.emplace_back([&, k, cmd] { rc[k] = run_on(user, hl[k], cmd, dry); });
    }
    for (auto &t : ths) t.join();
    int bad = 0;
    for (size_t k = 0; k < rc.size(); ++k) {
        if (rc[k]) {
            fprintf(stderr, "%s exited with %d\n", hl[k].public_addr.c_str(), rc[k]);
            bad = rc[k];
        }
    }
    return bad;
}
