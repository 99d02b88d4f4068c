// b200run - single-node job launcher for one-process-per-GPU programs on an NVSwitch box.
//
// Replaces, for this framework, the launchers the reference shells out to:
//   * `mpirun` / `mpiexec` (Open MPI and Intel MPI spellings used in launch_openmpi.sh:19-26,
//     launch_intelmpi.sh:24-30, collectives/launch_openmpi.sh:6-13, collectives/3d/launch_mpiccl.sh:24)
//       -np/-n N  -ppn N  -genv K=V | -genv K V  -x VAR[=V]  --bind-to core|socket|none
//       --map-by socket:PE=n | slot:PE=n | numa:PE=n  --report-bindings  --oversubscribe  --allow-run-as-root
//   * the `deepspeed` CLI (collectives/launch_dsccl.sh:7-15, collectives/3d/launch_dsccl.sh:69-74)
//       --master_port P  --master_addr A  --include localhost:0,1,...  --num_gpus N
//       --bind_cores_to_rank  --bind_core_list a,b-c,...  --no_local_rank   <script.py> [args]
// The personality is chosen from argv[0] (bin/mpirun and bin/deepspeed are symlinks to this binary)
// or with `--as mpirun|deepspeed`.
//
// What it does: discovers GPUs and their NUMA affinity (NVML, loaded at run time; absent on CPU-only
// machines), assigns rank i to GPU i, binds each rank to host cores (PE cores per rank, packed per
// NUMA node of its GPU when known), exports the environment both ecosystems expect (RANK,
// LOCAL_RANK, WORLD_SIZE, LOCAL_WORLD_SIZE, MASTER_ADDR, MASTER_PORT, OMPI_COMM_WORLD_*, PMI_*,
// DLBB_JOB_ID), puts the compat shims (mpi4py / deepspeed / oneccl_bindings_for_pytorch) on
// PYTHONPATH, spawns the ranks, and tears the whole job down with the first failing rank's exit
// code if any rank dies.
#include <dirent.h>
#include <dlfcn.h>
#include <errno.h>
#include <limits.h>
#include <sched.h>
#include <signal.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <netinet/in.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <map>
#include <set>
#include <string>
#include <vector>

namespace {

enum class Persona { Native, Mpirun, Deepspeed };

struct Options {
    Persona persona = Persona::Native;
    int nproc = -1;
    int ppn = -1;
    int pe = 0;                        // cores per rank (0 = no binding requested)
    std::string bind_to = "";          // core | socket | numa | none
    std::string map_by = "slot";       // slot | socket | numa
    bool report_bindings = false;
    bool tag_output = false;
    bool bind_cores_to_rank = false;   // deepspeed
    std::vector<std::vector<int>> bind_core_list;  // deepspeed: one entry per rank
    std::vector<int> include_gpus;     // deepspeed --include localhost:...
    bool no_local_rank = true;         // we never append --local_rank (the reference scripts take no args)
    std::string master_addr = "127.0.0.1";
    int master_port = 0;
    std::vector<std::pair<std::string, std::string>> env;  // forwarded / set for every rank
    std::vector<std::string> cmd;
    bool dry_run = false;
};

struct GpuInfo {
    int index;
    int numa = -1;
    std::vector<int> cpus;   // affinity mask from NVML (may be empty)
};

void die(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    fprintf(stderr, "b200run: ");
    vfprintf(stderr, fmt, ap);
    fprintf(stderr, "\n");
    va_end(ap);
    exit(2);
}

std::vector<int> parse_int_list(const std::string& s) {  // "0,2-4,7"
    std::vector<int> out;
    size_t i = 0;
    while (i < s.size()) {
        size_t j = s.find(',', i);
        if (j == std::string::npos) j = s.size();
        std::string tok = s.substr(i, j - i);
        if (!tok.empty()) {
            size_t d = tok.find('-');
            if (d != std::string::npos && d > 0) {
                int a = atoi(tok.substr(0, d).c_str()), b = atoi(tok.substr(d + 1).c_str());
                for (int v = a; v <= b; ++v) out.push_back(v);
            } else {
                out.push_back(atoi(tok.c_str()));
            }
        }
        i = j + 1;
    }
    return out;
}

std::string cpus_to_string(const std::vector<int>& c) {
    std::string s;
    for (size_t i = 0; i < c.size();) {
        size_t j = i;
        while (j + 1 < c.size() && c[j + 1] == c[j] + 1) ++j;
        if (!s.empty()) s += ",";
        s += std::to_string(c[i]);
        if (j > i) s += "-" + std::to_string(c[j]);
        i = j + 1;
    }
    return s.empty() ? "(unbound)" : s;
}

// ---------------------------------------------------------------- topology
std::vector<int> online_cpus() {
    cpu_set_t set;
    CPU_ZERO(&set);
    std::vector<int> out;
    if (sched_getaffinity(0, sizeof(set), &set) == 0)
        for (int i = 0; i < CPU_SETSIZE; ++i)
            if (CPU_ISSET(i, &set)) out.push_back(i);
    return out;
}

std::map<int, std::vector<int>> numa_nodes() {  // node -> cpus, from sysfs
    std::map<int, std::vector<int>> out;
    DIR* d = opendir("/sys/devices/system/node");
    if (!d) return out;
    while (dirent* e = readdir(d)) {
        if (strncmp(e->d_name, "node", 4) != 0 || !isdigit(e->d_name[4])) continue;
        const int node = atoi(e->d_name + 4);
        std::string path = std::string("/sys/devices/system/node/") + e->d_name + "/cpulist";
        FILE* f = fopen(path.c_str(), "r");
        if (!f) continue;
        char buf[4096] = {0};
        if (fgets(buf, sizeof(buf), f)) {
            std::string s(buf);
            while (!s.empty() && (s.back() == '\n' || s.back() == ' ')) s.pop_back();
            out[node] = parse_int_list(s);
        }
        fclose(f);
    }
    closedir(d);
    return out;
}

// NVML through dlopen: the launcher must also run where no driver is installed.
std::vector<GpuInfo> discover_gpus(bool verbose) {
    std::vector<GpuInfo> gpus;
    void* h = dlopen("libnvidia-ml.so.1", RTLD_NOW | RTLD_LOCAL);
    if (!h) return gpus;
    typedef int (*init_t)();
    typedef int (*count_t)(unsigned*);
    typedef int (*handle_t)(unsigned, void**);
    typedef int (*affin_t)(void*, unsigned, unsigned long*);
    typedef int (*nvlink_t)(void*, unsigned, unsigned*);
    auto init = reinterpret_cast<init_t>(dlsym(h, "nvmlInit_v2"));
    auto count = reinterpret_cast<count_t>(dlsym(h, "nvmlDeviceGetCount_v2"));
    auto handle = reinterpret_cast<handle_t>(dlsym(h, "nvmlDeviceGetHandleByIndex_v2"));
    auto affin = reinterpret_cast<affin_t>(dlsym(h, "nvmlDeviceGetCpuAffinity"));
    auto nvlink = reinterpret_cast<nvlink_t>(dlsym(h, "nvmlDeviceGetNvLinkState"));
    unsigned n = 0;
    if (!init || !count || !handle || init() != 0 || count(&n) != 0) return gpus;
    for (unsigned i = 0; i < n; ++i) {
        GpuInfo g;
        g.index = static_cast<int>(i);
        void* dev = nullptr;
        if (handle(i, &dev) == 0 && dev) {
            unsigned long mask[16] = {0};
            if (affin && affin(dev, 16, mask) == 0)
                for (int w = 0; w < 16; ++w)
                    for (int b = 0; b < 64; ++b)
                        if (mask[w] & (1ul << b)) g.cpus.push_back(w * 64 + b);
            if (verbose && nvlink) {
                int up = 0;
                for (unsigned l = 0; l < 18; ++l) {
                    unsigned st = 0;
                    if (nvlink(dev, l, &st) == 0 && st) ++up;
                }
                fprintf(stderr, "b200run: GPU %u: %d NVLink links up, %zu affine host cores\n", i, up, g.cpus.size());
            }
        }
        gpus.push_back(g);
    }
    return gpus;
}

int free_tcp_port() {
    int s = socket(AF_INET, SOCK_STREAM, 0);
    if (s < 0) return 29500;
    sockaddr_in a{};
    a.sin_family = AF_INET;
    a.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
    a.sin_port = 0;
    int port = 29500;
    socklen_t len = sizeof(a);
    if (bind(s, reinterpret_cast<sockaddr*>(&a), sizeof(a)) == 0 && getsockname(s, reinterpret_cast<sockaddr*>(&a), &len) == 0)
        port = ntohs(a.sin_port);
    close(s);
    return port;
}

// ---------------------------------------------------------------- argument parsing
void add_env_spec(Options& o, const std::string& spec, const char* next, int& i) {
    // "K=V" sets; bare "K" forwards the launcher's own value (Open MPI -x) or takes the next arg (Intel -genv K V)
    size_t eq = spec.find('=');
    if (eq != std::string::npos) {
        o.env.emplace_back(spec.substr(0, eq), spec.substr(eq + 1));
    } else if (next != nullptr) {
        o.env.emplace_back(spec, next);
        ++i;
    } else if (const char* v = getenv(spec.c_str())) {
        o.env.emplace_back(spec, v);
    }
}

void parse_map_by(Options& o, const std::string& v) {
    // socket:PE=14, slot:PE=2, numa:PE=4, ppr:1:socket:PE=4 (last object wins)
    std::string s = v;
    size_t pe = s.find("PE=");
    if (pe != std::string::npos) o.pe = atoi(s.c_str() + pe + 3);
    if (s.find("socket") != std::string::npos) o.map_by = "socket";
    else if (s.find("numa") != std::string::npos) o.map_by = "numa";
    else o.map_by = "slot";
}

Options parse(int argc, char** argv) {
    Options o;
    std::string base = argv[0];
    size_t sl = base.rfind('/');
    if (sl != std::string::npos) base = base.substr(sl + 1);
    if (base == "mpirun" || base == "mpiexec" || base == "mpiexec.hydra") o.persona = Persona::Mpirun;
    else if (base == "deepspeed" || base == "ds") o.persona = Persona::Deepspeed;
    int i = 1;
    for (; i < argc; ++i) {
        std::string a = argv[i];
        auto need = [&](const char* what) -> std::string {
            if (i + 1 >= argc) die("%s needs a value", what);
            return argv[++i];
        };
        if (a == "--as") {
            std::string p = need("--as");
            o.persona = p == "mpirun" ? Persona::Mpirun : p == "deepspeed" ? Persona::Deepspeed : Persona::Native;
        } else if (a == "-np" || a == "-n" || a == "--np" || a == "--nproc-per-node" || a == "--nproc_per_node" || a == "-c") {
            o.nproc = atoi(need("-np").c_str());
        } else if (a == "-ppn" || a == "--ppn" || a == "-perhost") {
            o.ppn = atoi(need("-ppn").c_str());
        } else if (a == "-genv" || a == "-env") {
            std::string spec = need("-genv");
            const char* nxt = (spec.find('=') == std::string::npos && i + 1 < argc) ? argv[i + 1] : nullptr;
            add_env_spec(o, spec, nxt, i);
        } else if (a == "-x") {
            std::string spec = need("-x");
            add_env_spec(o, spec, nullptr, i);
        } else if (a == "--bind-to" || a == "-bind-to") {
            o.bind_to = need("--bind-to");
        } else if (a == "--map-by" || a == "-map-by") {
            parse_map_by(o, need("--map-by"));
        } else if (a == "--report-bindings" || a == "-report-bindings") {
            o.report_bindings = true;
        } else if (a == "--tag-output" || a == "-l" || a == "-prepend-rank") {
            o.tag_output = true;
        } else if (a == "--oversubscribe" || a == "--allow-run-as-root" || a == "--use-hwthread-cpus") {
            // accepted for compatibility; nothing to do on a single node
        } else if (a == "--master_port" || a == "--master-port") {
            o.master_port = atoi(need("--master_port").c_str());
        } else if (a == "--master_addr" || a == "--master-addr") {
            o.master_addr = need("--master_addr");
        } else if (a == "--include") {
            std::string v = need("--include");      // localhost:0,1,2,3
            size_t c = v.find(':');
            o.include_gpus = parse_int_list(c == std::string::npos ? v : v.substr(c + 1));
            o.nproc = static_cast<int>(o.include_gpus.size());
        } else if (a == "--num_gpus" || a == "--num-gpus" || a == "--num_accelerators") {
            o.nproc = atoi(need("--num_gpus").c_str());
        } else if (a == "--num_nodes" || a == "--hostfile" || a == "-H" || a == "--host" || a == "-hosts") {
            need(a.c_str());  // single node only: value ignored
        } else if (a == "--bind_cores_to_rank") {
            o.bind_cores_to_rank = true;
        } else if (a == "--bind_core_list") {
            // one comma-separated entry per rank; an entry may itself be a range "a-b"
            std::string v = need("--bind_core_list");
            size_t p = 0;
            while (p <= v.size()) {
                size_t q = v.find(',', p);
                if (q == std::string::npos) q = v.size();
                if (q > p) o.bind_core_list.push_back(parse_int_list(v.substr(p, q - p)));
                p = q + 1;
            }
        } else if (a == "--no_local_rank") {
            o.no_local_rank = true;
        } else if (a == "--dry-run") {
            o.dry_run = true;
        } else if (a == "--") {
            ++i;
            break;
        } else if (!a.empty() && a[0] == '-' && o.cmd.empty()) {
            die("unknown option %s", a.c_str());
        } else {
            break;
        }
    }
    for (; i < argc; ++i) o.cmd.push_back(argv[i]);
    if (o.cmd.empty()) die("no program given");
    // `deepspeed script.py` runs the script with the Python interpreter
    if (o.persona == Persona::Deepspeed) {
        const std::string& s = o.cmd[0];
        if (s.size() > 3 && s.substr(s.size() - 3) == ".py") {
            const char* py = getenv("DLBB_PYTHON");
            o.cmd.insert(o.cmd.begin(), {py ? py : "python", "-u"});
        }
    }
    return o;
}

// ---------------------------------------------------------------- binding plan
std::vector<std::vector<int>> plan_bindings(const Options& o, int nproc, const std::vector<GpuInfo>& gpus) {
    std::vector<std::vector<int>> plan(nproc);
    if (!o.bind_core_list.empty()) {  // deepspeed: explicit per-rank list
        for (int r = 0; r < nproc && r < (int)o.bind_core_list.size(); ++r) plan[r] = o.bind_core_list[r];
        return plan;
    }
    const std::vector<int> avail = online_cpus();
    if (avail.empty() || o.bind_to == "none") return plan;
    int pe = o.pe;
    if (pe <= 0 && (o.bind_cores_to_rank || o.bind_to == "core")) pe = std::max<int>(1, (int)avail.size() / std::max(1, nproc));
    if (pe <= 0) return plan;
    const auto nodes = numa_nodes();
    const std::set<int> avail_set(avail.begin(), avail.end());
    if ((o.map_by == "socket" || o.map_by == "numa") && nodes.size() > 1) {
        // round-robin ranks over NUMA nodes, PE cores each, packed inside the node
        std::vector<std::vector<int>> pools;
        for (auto& kv : nodes) {
            std::vector<int> p;
            for (int c : kv.second) if (avail_set.count(c)) p.push_back(c);
            if (!p.empty()) pools.push_back(p);
        }
        std::vector<size_t> cursor(pools.size(), 0);
        for (int r = 0; r < nproc && !pools.empty(); ++r) {
            size_t n = r % pools.size();
            // prefer the NUMA node of the rank's GPU when known
            if (r < (int)gpus.size() && !gpus[r].cpus.empty())
                for (size_t k = 0; k < pools.size(); ++k)
                    if (std::find(pools[k].begin(), pools[k].end(), gpus[r].cpus[0]) != pools[k].end()) { n = k; break; }
            for (int k = 0; k < pe; ++k) {
                if (cursor[n] >= pools[n].size()) cursor[n] = 0;  // oversubscribed: wrap
                plan[r].push_back(pools[n][cursor[n]++]);
            }
        }
        return plan;
    }
    // slot mapping: consecutive blocks of PE cores
    size_t cur = 0;
    for (int r = 0; r < nproc; ++r)
        for (int k = 0; k < pe; ++k) {
            if (cur >= avail.size()) cur = 0;
            plan[r].push_back(avail[cur++]);
        }
    return plan;
}

std::string self_dir() {
    char buf[PATH_MAX];
    ssize_t n = readlink("/proc/self/exe", buf, sizeof(buf) - 1);
    if (n <= 0) return ".";
    buf[n] = 0;
    std::string p(buf);
    size_t sl = p.rfind('/');
    return sl == std::string::npos ? "." : p.substr(0, sl);
}

volatile sig_atomic_t g_signal = 0;
void on_signal(int s) { g_signal = s; }

}  // namespace

int main(int argc, char** argv) {
    Options o = parse(argc, argv);
    const bool verbose = getenv("DLBB_LOG_LEVEL") != nullptr;
    std::vector<GpuInfo> gpus = discover_gpus(verbose || o.report_bindings);
    int nproc = o.nproc;
    if (nproc <= 0) nproc = gpus.empty() ? 1 : static_cast<int>(gpus.size());
    if (o.ppn > 0 && o.ppn < nproc) die("-ppn %d < -np %d: multi-node launch is not supported (single NVSwitch node)", o.ppn, nproc);
    if (!gpus.empty() && nproc > (int)gpus.size() && getenv("DLBB_ALLOW_GPU_OVERSUBSCRIBE") == nullptr && verbose)
        fprintf(stderr, "b200run: %d ranks on %zu GPUs: ranks share devices\n", nproc, gpus.size());
    if (o.master_port == 0) {
        const char* mp = getenv("MASTER_PORT");
        o.master_port = mp ? atoi(mp) : free_tcp_port();
    }
    const std::vector<std::vector<int>> plan = plan_bindings(o, nproc, gpus);
    // environment common to all ranks
    const std::string root = self_dir() + "/..";
    std::string pypath = root + "/compat:" + root;
    if (const char* p = getenv("PYTHONPATH")) pypath += std::string(":") + p;
    char job[64];
    snprintf(job, sizeof(job), "%d-%ld", o.master_port, (long)getpid());

    if (o.report_bindings || o.dry_run)
        for (int r = 0; r < nproc; ++r) {
            const int gpu = !o.include_gpus.empty() ? o.include_gpus[r] : (gpus.empty() ? -1 : r % (int)gpus.size());
            fprintf(stderr, "[b200run] rank %d -> gpu %d, bound to cores %s\n", r, gpu, cpus_to_string(plan[r]).c_str());
        }
    if (o.dry_run) return 0;

    struct sigaction sa{};
    sa.sa_handler = on_signal;
    sigaction(SIGINT, &sa, nullptr);
    sigaction(SIGTERM, &sa, nullptr);

    std::vector<pid_t> pids(nproc, -1);
    for (int r = 0; r < nproc; ++r) {
        int pipefd[2] = {-1, -1};
        if (o.tag_output && pipe(pipefd) != 0) die("pipe: %s", strerror(errno));
        pid_t pid = fork();
        if (pid < 0) die("fork: %s", strerror(errno));
        if (pid == 0) {
            setpgid(0, 0);
            if (!plan[r].empty()) {
                cpu_set_t set;
                CPU_ZERO(&set);
                for (int c : plan[r]) if (c >= 0 && c < CPU_SETSIZE) CPU_SET(c, &set);
                sched_setaffinity(0, sizeof(set), &set);
            }
            auto set = [](const char* k, const std::string& v) { setenv(k, v.c_str(), 1); };
            const std::string R = std::to_string(r), N = std::to_string(nproc);
            set("RANK", R); set("LOCAL_RANK", R); set("WORLD_SIZE", N); set("LOCAL_WORLD_SIZE", N);
            set("MASTER_ADDR", o.master_addr); set("MASTER_PORT", std::to_string(o.master_port));
            set("OMPI_COMM_WORLD_RANK", R); set("OMPI_COMM_WORLD_SIZE", N); set("OMPI_COMM_WORLD_LOCAL_RANK", R);
            set("OMPI_COMM_WORLD_LOCAL_SIZE", N);
            set("PMI_RANK", R); set("PMI_SIZE", N); set("MPI_LOCALRANKID", R); set("MPI_LOCALNRANKS", N);
            set("CROSS_RANK", "0"); set("CROSS_SIZE", "1");
            set("DLBB_JOB_ID", job);
            set("DLBB_LAUNCHER", o.persona == Persona::Mpirun ? "mpirun" : o.persona == Persona::Deepspeed ? "deepspeed" : "b200run");
            set("PYTHONPATH", pypath);
            if (!o.include_gpus.empty()) {
                std::string vis;
                for (size_t k = 0; k < o.include_gpus.size(); ++k) vis += (k ? "," : "") + std::to_string(o.include_gpus[k]);
                set("CUDA_VISIBLE_DEVICES", vis);
            }
            if (!plan[r].empty() && getenv("OMP_NUM_THREADS") == nullptr) set("OMP_NUM_THREADS", std::to_string(plan[r].size()));
            for (auto& kv : o.env) setenv(kv.first.c_str(), kv.second.c_str(), 1);
            if (o.tag_output) {
                close(pipefd[0]);
                dup2(pipefd[1], 1);
                dup2(pipefd[1], 2);
                close(pipefd[1]);
            }
            std::vector<char*> av;
            for (auto& s : o.cmd) av.push_back(const_cast<char*>(s.c_str()));
            av.push_back(nullptr);
            execvp(av[0], av.data());
            fprintf(stderr, "b200run: cannot exec %s: %s\n", av[0], strerror(errno));
            _exit(127);
        }
        pids[r] = pid;
        if (o.tag_output) {
            close(pipefd[1]);
            // a tiny forwarder per rank that prefixes lines with the rank
            pid_t fw = fork();
            if (fw == 0) {
                FILE* in = fdopen(pipefd[0], "r");
                char* line = nullptr;
                size_t cap = 0;
                while (in && getline(&line, &cap, in) > 0) {
                    fprintf(stdout, "[%d] %s", r, line);
                    fflush(stdout);
                }
                _exit(0);
            }
            close(pipefd[0]);
        }
    }

    // wait: first failure (or a signal to the launcher) takes the whole job down
    int exit_code = 0, alive = nproc;
    bool killing = false;
    time_t kill_deadline = 0;
    while (alive > 0) {
        int status = 0;
        pid_t p = waitpid(-1, &status, WNOHANG);
        if (p > 0) {
            auto it = std::find(pids.begin(), pids.end(), p);
            if (it == pids.end()) continue;  // an output forwarder
            const int r = static_cast<int>(it - pids.begin());
            *it = -1;
            --alive;
            int code = 0;
            if (WIFEXITED(status)) code = WEXITSTATUS(status);
            else if (WIFSIGNALED(status)) code = 128 + WTERMSIG(status);
            if (code != 0 && exit_code == 0) {
                exit_code = code;
                if (!killing) fprintf(stderr, "b200run: rank %d exited with code %d; terminating the job\n", r, code);
            }
            if (code != 0 && !killing) {
                killing = true;
                kill_deadline = time(nullptr) + 5;
                for (pid_t q : pids) if (q > 0) kill(-q, SIGTERM);
            }
            continue;
        }
        if (g_signal && !killing) {
            killing = true;
            kill_deadline = time(nullptr) + 5;
            if (exit_code == 0) exit_code = 128 + g_signal;
            for (pid_t q : pids) if (q > 0) kill(-q, g_signal);
        }
        if (killing && time(nullptr) > kill_deadline) {
            for (pid_t q : pids) if (q > 0) kill(-q, SIGKILL);
            kill_deadline = time(nullptr) + 3600;
        }
        usleep(20000);
    }
    return exit_code;
}
