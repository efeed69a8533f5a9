// cellector — host CLI over libcellector_cuda.so: the reference's command line, stdout and output files
// (cellector/src/params.yml, cellector/src/main.rs:25-50) with the scoring EM on the GPU behind the C ABI of
// include/cellector_cuda.h.  A Rust host would make the same calls through the extern "C" block of
// INTEGRATION.md; no Rust toolchain exists in this image, so the host is C++.
//
//   cellector -a alt.mtx -r ref.mtx -b barcodes.tsv --output_directory OUT [--min_alt 4] [--min_ref 4]
//             [-g ground_truth.tsv] [-v variants.vcf] [--posterior_threshold 0.999]
//             [--interquartile_range_multiple 5] [--min_alleles_posterior 5] [--expected_percent_minority P]
//             [--min_loci_for_assignment 30]
//   additive flags (not in the reference): --device N, --threads N, --max_iterations N, --timing_json FILE,
//             --no_iteration_files (skip the per-iteration diagnostics), --parse_only (parser self-check),
//             --partition_only (prints the --gpus N cell ranges, no CUDA),
//             --gpus N: cells sharded over N GPUs (devices --device .. --device + N - 1).  The process parses the
//             matrix once, then forks N - 1 workers BEFORE any CUDA call; every process owns one GPU and one context
//             (the library's model: one process per GPU, clx_comm_init between them), takes a contiguous range of
//             cells with an equal share of the nonzeros, and rank 0 writes every file from per-cell arrays the ranks
//             fill in shared memory.
//
// There is no CPU fallback: without a CUDA device the run fails in clx_create.
#include <sys/mman.h>
#include <sys/prctl.h>
#include <sys/stat.h>
#include <sys/wait.h>
#include <signal.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <numeric>
#include <stdexcept>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../../include/cellector_cuda.h"
#include "mtx_parse.h"
#include "rust_fmt.h"

using namespace clxhost;

namespace {

struct Params {  // main.rs:613-627
    std::string ref_mtx, alt_mtx, barcodes, output_directory;
    uint64_t min_alt = 4, min_ref = 4;
    bool has_ground_truth = false, has_vcf = false;
    std::string ground_truth, vcf;
    double posterior_threshold = 0.999;
    double interquartile_range_multiple = 5.0;
    uint64_t min_alleles_posterior = 5;  // parsed, changes nothing (main.rs:301-303)
    bool has_expected_percent_minority = false;
    double expected_percent_minority = 0.0;  // parsed, never used (main.rs:654-657)
    uint64_t min_loci_used = 30;
    // additive
    int device = 0;
    int gpus = 1;
    unsigned threads = 0;
    uint64_t max_iterations = 0;  // 0 = no cap, like the reference (main.rs:42-46)
    std::string timing_json;
    bool no_iteration_files = false, parse_only = false, partition_only = false;
};

// --gpus N: what the ranks share (one anonymous MAP_SHARED mapping made before the fork)
struct TeamShared {
    std::atomic<uint32_t> arrived, generation;  // barrier
    std::atomic<uint32_t> id_ready;             // rank 0 has written nccl_id
    std::atomic<uint32_t> failed;               // a rank died: everybody leaves
    unsigned char nccl_id[128];
};
TeamShared* g_team = nullptr;
int g_rank = 0, g_world = 1;
std::vector<pid_t> g_children;

[[noreturn]] void die(const std::string& msg, int code = 101) {  // 101 = exit status of a Rust panic
    if (g_world > 1) fprintf(stderr, "cellector (rank %d): %s\n", g_rank, msg.c_str());
    else fprintf(stderr, "cellector: %s\n", msg.c_str());
    if (g_team) g_team->failed.store(1);
    if (g_rank != 0) _exit(code);  // a worker must not flush stdio buffers inherited from the parent
    for (const pid_t pid : g_children) kill(pid, SIGKILL);
    exit(code);
}

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

uint64_t parse_u64_arg(const std::string& name, const std::string& v) {
    const char* p = v.data();
    uint64_t out = 0;
    if (!next_usize(p, v.data() + v.size(), &out) || p != v.data() + v.size()) die("invalid value '" + v + "' for --" + name);
    return out;
}
double parse_f64_arg(const std::string& name, const std::string& v) {
    char* end = nullptr;
    const double d = strtod(v.c_str(), &end);
    if (v.empty() || end != v.c_str() + v.size()) die("invalid value '" + v + "' for --" + name);
    return d;
}

void usage() {
    fputs(
        "cellector 1.0.0 (CUDA sm_100a build)\ngenotype outlier detection for scRNAseq\n\nUSAGE:\n"
        "    cellector [OPTIONS] --alt <alt> --barcodes <barcodes> --output_directory <output_directory> --ref <ref>\n\n"
        "OPTIONS:\n"
        "    -a, --alt <alt>                          alt.mtx matrix from vartrix\n"
        "    -b, --barcodes <barcodes>                cell barcodes\n"
        "    -r, --ref <ref>                          ref.mtx matrix from vartrix\n"
        "        --output_directory <dir>             output directory where other output files will be stored\n"
        "    -g, --ground_truth <file>                barcode<TAB>assignment ground truth\n"
        "    -v, --vcf <file>                         vcf associated with alt.mtx and ref.mtx\n"
        "        --min_alt <n>                        minimum number of cells containing the alt allele (default 4)\n"
        "        --min_ref <n>                        minimum number of cells containing the ref allele (default 4)\n"
        "        --posterior_threshold <p>            posterior threshold for assignment (default 0.999)\n"
        "        --interquartile_range_multiple <m>   IQR multiples below the 25th percentile for the outlier threshold (default 5)\n"
        "        --min_alleles_posterior <n>          (default 5)\n"
        "        --expected_percent_minority <p>\n"
        "        --min_loci_for_assignment <n>        (default 30)\n"
        "        --device <n> --gpus <n> --threads <n> --max_iterations <n> --timing_json <file> --no_iteration_files --parse_only\n",
        stderr);
}

Params load_params(int argc, char** argv) {
    static const std::map<std::string, std::string> shorts = {{"-r", "ref"}, {"-a", "alt"}, {"-b", "barcodes"}, {"-g", "ground_truth"}, {"-v", "vcf"}};
    std::map<std::string, std::string> kv;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i], name, val;
        bool have_val = false;
        if (a == "-h" || a == "--help") {
            usage();
            exit(0);
        }
        if (a.rfind("--", 0) == 0) {
            const size_t eq = a.find('=');
            name = a.substr(2, eq == std::string::npos ? std::string::npos : eq - 2);
            if (eq != std::string::npos) {
                val = a.substr(eq + 1);
                have_val = true;
            }
        } else if (a.size() >= 2 && a[0] == '-' && shorts.count(a.substr(0, 2))) {
            name = shorts.at(a.substr(0, 2));
            if (a.size() > 2) {  // -afile or -a=file
                val = a.substr(a[2] == '=' ? 3 : 2);
                have_val = true;
            }
        } else {
            usage();
            die("error: Found argument '" + a + "' which wasn't expected, or isn't valid in this context", 1);
        }
        if (name == "no_iteration_files" || name == "parse_only" || name == "partition_only") {
            kv[name] = "1";
            continue;
        }
        if (!have_val) {
            if (i + 1 >= argc) die("error: The argument '--" + name + "' requires a value but none was supplied", 1);
            val = argv[++i];
        }
        kv[name] = val;
    }
    static const char* known[] = {"output_directory", "ref", "alt", "barcodes", "min_alt", "min_ref", "ground_truth", "vcf",
                                  "posterior_threshold", "interquartile_range_multiple", "min_alleles_posterior",
                                  "expected_percent_minority", "min_loci_for_assignment", "device", "gpus", "threads", "max_iterations",
                                  "timing_json", "no_iteration_files", "parse_only", "partition_only"};
    for (const auto& e : kv) {
        bool ok = false;
        for (const char* k : known) ok = ok || e.first == k;
        if (!ok) {
            usage();
            die("error: Found argument '--" + e.first + "' which wasn't expected, or isn't valid in this context", 1);
        }
    }
    Params p;
    for (const char* req : {"output_directory", "ref", "alt", "barcodes"})
        if (!kv.count(req)) {
            usage();
            die(std::string("error: The following required arguments were not provided:\n    --") + req, 1);
        }
    p.output_directory = kv["output_directory"];
    p.ref_mtx = kv["ref"];
    p.alt_mtx = kv["alt"];
    p.barcodes = kv["barcodes"];
    if (kv.count("min_alt")) p.min_alt = parse_u64_arg("min_alt", kv["min_alt"]);
    if (kv.count("min_ref")) p.min_ref = parse_u64_arg("min_ref", kv["min_ref"]);
    if (kv.count("ground_truth")) { p.has_ground_truth = true; p.ground_truth = kv["ground_truth"]; }
    if (kv.count("vcf")) { p.has_vcf = true; p.vcf = kv["vcf"]; }
    if (kv.count("posterior_threshold")) p.posterior_threshold = parse_f64_arg("posterior_threshold", kv["posterior_threshold"]);
    if (kv.count("interquartile_range_multiple"))
        p.interquartile_range_multiple = parse_f64_arg("interquartile_range_multiple", kv["interquartile_range_multiple"]);
    if (kv.count("min_alleles_posterior")) p.min_alleles_posterior = parse_u64_arg("min_alleles_posterior", kv["min_alleles_posterior"]);
    if (kv.count("expected_percent_minority")) {
        p.has_expected_percent_minority = true;
        p.expected_percent_minority = parse_f64_arg("expected_percent_minority", kv["expected_percent_minority"]);
    }
    if (kv.count("min_loci_for_assignment")) p.min_loci_used = parse_u64_arg("min_loci_for_assignment", kv["min_loci_for_assignment"]);
    if (kv.count("device")) p.device = (int)parse_u64_arg("device", kv["device"]);
    if (kv.count("gpus")) p.gpus = (int)parse_u64_arg("gpus", kv["gpus"]);
    if (p.gpus < 1 || p.gpus > 64) die("--gpus must be between 1 and 64", 1);
    if (kv.count("threads")) p.threads = (unsigned)parse_u64_arg("threads", kv["threads"]);
    if (kv.count("max_iterations")) p.max_iterations = parse_u64_arg("max_iterations", kv["max_iterations"]);
    if (kv.count("timing_json")) p.timing_json = kv["timing_json"];
    p.no_iteration_files = kv.count("no_iteration_files") > 0;
    p.parse_only = kv.count("parse_only") > 0;
    p.partition_only = kv.count("partition_only") > 0;
    if (p.threads == 0) p.threads = std::max(1u, std::thread::hardware_concurrency());
    return p;
}

// load_data::create_output_dir (load_data.rs:66-71): spawns `mkdir DIR` (no -p) and ignores its failure
void create_output_dir(const Params& p) {
    const pid_t pid = fork();
    if (pid == 0) {
        FILE* devnull = freopen("/dev/null", "w", stderr);
        (void)devnull;
        execlp("mkdir", "mkdir", p.output_directory.c_str(), (char*)nullptr);
        _exit(127);
    }
    if (pid > 0) {
        int st = 0;
        waitpid(pid, &st, 0);
    }
}

std::vector<std::string> read_lines(const std::string& path) {
    FileBytes f(path);
    std::vector<std::string> out;
    for_each_line(f.data(), f.size(), [&](const char* s, size_t n) { out.emplace_back(s, n); });
    return out;
}

// load_data::load_barcodes (load_data.rs:73-83): line i -> cell id i; a repeated barcode keeps its LAST index
void load_barcodes(const Params& p, std::vector<std::string>* barcodes, std::unordered_map<std::string, size_t>* ids) {
    *barcodes = read_lines(p.barcodes);
    for (size_t i = 0; i < barcodes->size(); ++i) (*ids)[(*barcodes)[i]] = i;
}

// load_data::load_ground_truth (load_data.rs:85-107): one entry per DISTINCT barcode, default "na"
std::vector<std::string> load_ground_truth(const Params& p, const std::unordered_map<std::string, size_t>& ids) {
    std::vector<std::string> gt(ids.size(), "na");
    if (!p.has_ground_truth) return gt;
    for (const std::string& line : read_lines(p.ground_truth)) {
        const size_t tab = line.find('\t');
        if (tab == std::string::npos || line.find('\t', tab + 1) != std::string::npos)
            die("Invalid line format: " + line + "\nThe correct format is: barcode\tassignment");
        const auto it = ids.find(line.substr(0, tab));
        if (it != ids.end() && it->second < gt.size()) gt[it->second] = line.substr(tab + 1);
    }
    return gt;
}

struct VcfLocus {  // load_data.rs:28-35
    std::string chrom, pos;
};
std::vector<VcfLocus> load_vcf_data(const Params& p) {
    std::vector<VcfLocus> out;
    for (const std::string& line : read_lines(p.vcf)) {
        if (!line.empty() && line[0] == '#') continue;
        std::vector<std::string> toks;
        size_t i = 0;
        while (toks.size() < 5) {
            const size_t t = line.find('\t', i);
            toks.push_back(line.substr(i, t == std::string::npos ? std::string::npos : t - i));
            if (t == std::string::npos) break;
            i = t + 1;
        }
        if (toks.size() < 5) die("vcf record with fewer than 5 columns: " + line);
        out.push_back({toks[0], toks[1]});
    }
    return out;
}

struct FileWriter {  // BufWriter<File>
    FILE* f;
    std::string buf;
    explicit FileWriter(const std::string& path) : f(fopen(path.c_str(), "wb")) {
        if (!f) die("Unable to create file " + path);
        buf.reserve(1 << 20);
    }
    void flush() {
        if (!buf.empty()) fwrite(buf.data(), 1, buf.size(), f);
        buf.clear();
    }
    void maybe_flush() {
        if (buf.size() > (1 << 20) - 4096) flush();
    }
    ~FileWriter() {
        flush();
        fclose(f);
    }
};

#define CLX(call)                                                                                  \
    do {                                                                                           \
        const int rc__ = (call);                                                                   \
        if (rc__ != CLX_OK) die(std::string(#call) + " failed: " + clx_last_error(ctx));           \
    } while (0)

// statrs 0.16.0 pieces of the VCF writer (main.rs:92-97): Binomial::pmf via ln_binomial / ln_factorial
// (a 171-entry table of ln(i!) from the running f64 product; Lanczos ln_gamma beyond it)
double host_ln_gamma(double x) {
    static const double dk[11] = {2.48574089138753565546e-5, 1.05142378581721974210, -3.45687097222016235469,
                                  4.51227709466894823700, -2.98285225323576655721, 1.05639711577126713077,
                                  -1.95428773191645869583e-1, 1.70970543404441224307e-2, -5.71926117404305781283e-4,
                                  4.63399473359905636708e-6, -2.71994908488607703910e-9};
    const double r = 10.900511, ln_2_sqrt_e_over_pi = 0.6207822376352452223455184457816472122518527279025978;
    double s = dk[0];
    for (int i = 1; i < 11; ++i) s += dk[i] / (x + (double)i - 1.0);
    return log(s) + ln_2_sqrt_e_over_pi + (x - 0.5) * log((x - 0.5 + r) / 2.71828182845904523536028747135266250);
}
double host_ln_factorial(uint64_t n) {
    static double table[171];
    static bool init = false;
    if (!init) {
        double f = 1.0;
        table[0] = log(1.0);
        for (int i = 1; i <= 170; ++i) {
            f *= (double)i;
            table[i] = log(f);
        }
        init = true;
    }
    return n <= 170 ? table[n] : host_ln_gamma((double)n + 1.0);
}
double binomial_pmf(double p, uint64_t n, uint64_t k) {
    if (k > n) return 0.0;
    if (p == 0.0) return k == 0 ? 1.0 : 0.0;
    if (p == 1.0) return k == n ? 1.0 : 0.0;
    const double lb = host_ln_factorial(n) - host_ln_factorial(k) - host_ln_factorial(n - k);
    return exp(lb + (double)k * log(p) + (double)(n - k) * log(1.0 - p));
}

// all ranks arrive, or the run ends: a rank that died has set `failed`; rank 0 also notices a worker that crashed
void team_barrier() {
    if (g_world == 1) return;
    const uint32_t gen = g_team->generation.load();
    if (g_team->arrived.fetch_add(1) + 1 == (uint32_t)g_world) {
        g_team->arrived.store(0);
        g_team->generation.fetch_add(1);
        return;
    }
    for (uint64_t spins = 0; g_team->generation.load() == gen; ++spins) {
        if (g_team->failed.load()) die("another rank failed");
        if (g_rank == 0 && (spins & 255) == 255) {
            int st = 0;
            const pid_t pid = waitpid(-1, &st, WNOHANG);
            if (pid > 0 && !(WIFEXITED(st) && WEXITSTATUS(st) == 0)) die("a worker process ended abnormally");
        }
        usleep(20);
    }
}

void* shared_alloc(size_t bytes) {
    void* p = mmap(nullptr, bytes ? bytes : 1, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (p == MAP_FAILED) die("mmap of the shared result arrays failed");
    return p;
}

}  // namespace

int run(const Params& params);

int main(int argc, char** argv) {
    if (argc == 2 && std::string(argv[1]) == "--format_selftest") {  // test hook: hex f64 bit patterns on stdin -> Display strings
        char line[64];
        while (fgets(line, sizeof(line), stdin)) {
            const unsigned long long bits = strtoull(line, nullptr, 16);
            double v;
            memcpy(&v, &bits, 8);
            printf("%s\n", f64_str(v).c_str());
        }
        return 0;
    }
    const Params params = load_params(argc, argv);
    try {
        return run(params);
    } catch (const std::exception& e) {
        die(e.what());
    }
}

namespace {

uint64_t fnv1a(const uint32_t* a, uint64_t n) {
    uint64_t h = 1469598103934665603ull;
    for (uint64_t i = 0; i < n; ++i) {
        h ^= a[i];
        h *= 1099511628211ull;
    }
    return h;
}

// statrs Data::median over a copy: N odd -> x_(N/2+1); N even -> (x_(N/2) + x_(N/2+1)) / 2; empty -> NaN
double median_of(std::vector<double> v) {
    if (v.empty()) return NAN;
    const size_t k = v.size() / 2;
    std::nth_element(v.begin(), v.begin() + k, v.end());
    const double hi = v[k];
    if (v.size() % 2 == 1) return hi;
    const double lo = *std::max_element(v.begin(), v.begin() + k);
    return (lo + hi) / 2.0;
}

// main.rs:177-226
void pretty_print(const std::map<std::string, std::map<std::string, uint64_t>>& assignment_gt_counts,
                  const std::map<std::string, uint64_t>& gt_counts) {
    std::vector<std::pair<std::string, uint64_t>> count_vec(gt_counts.begin(), gt_counts.end());
    std::stable_sort(count_vec.begin(), count_vec.end(), [](const auto& a, const auto& b) { return a.second > b.second; });
    std::string sb;
    const std::string first_header = "cellector assignment   ";
    const std::string header = "      0      1      unassigned\n";
    sb += first_header + header;
    size_t xoffset = std::max<size_t>(3, first_header.size() + 2);
    sb += "cell_hashing";
    sb.append(xoffset > 12 ? xoffset - 12 : 0, ' ');
    sb += "|";
    sb.append(header.size() - 1, '-');
    sb += "|\n";
    auto lookup = [&](const char* a, const std::string& gt) -> uint64_t {
        const auto it = assignment_gt_counts.find(a);
        if (it == assignment_gt_counts.end()) return 0;
        const auto jt = it->second.find(gt);
        return jt == it->second.end() ? 0 : jt->second;
    };
    for (const auto& gc : count_vec) {
        const std::string& gt = gc.first;
        xoffset = std::max(xoffset, gt.size() + 3);
        const std::string c0 = std::to_string(lookup("0", gt)), c1 = std::to_string(lookup("1", gt)),
                          un = std::to_string(lookup("unassigned", gt));
        sb += gt;
        const size_t gl = gt.size() ? gt.size() - 1 : 0;
        sb.append(xoffset > gl ? xoffset - gl : 0, ' ');
        sb += " |  " + c0;
        sb.append(c0.size() < 4 ? 4 - c0.size() : 0, ' ');
        sb += " |  " + c1;
        sb.append(c1.size() < 4 ? 4 - c1.size() : 0, ' ');
        sb += " |  " + un;
        sb.append(un.size() < 12 ? 12 - un.size() : 0, ' ');
        sb += "|\n";
    }
    sb.append(xoffset, ' ');
    sb += "|";
    sb.append(header.size() - 1, '-');
    sb += "|\n";
    printf("\n\n%s\n", sb.c_str());
}

}  // namespace

int run(const Params& params) {
    const double t_start = now_s();
    create_output_dir(params);
    std::vector<std::string> barcodes;
    std::unordered_map<std::string, size_t> barcode_ids;
    load_barcodes(params, &barcodes, &barcode_ids);
    const std::vector<std::string> ground_truth = load_ground_truth(params, barcode_ids);

    // ---- one parse of the alt/ref pair into pinned COO (load_data.rs:134-181 read them twice)
    const double t_parse0 = now_s();
    FileBytes alt(params.alt_mtx), ref(params.ref_mtx);
    const MtxPairPlan plan = plan_mtx_pair(alt, ref, params.threads);
    const uint64_t nnz0 = plan.n_pairs;
    if (plan.hdr.n_loci > 0xFFFFFFFEull || plan.hdr.n_cells > 0xFFFFFFFEull) die("matrix dimensions exceed 2^32 - 2");
    const uint32_t n_loci0 = (uint32_t)plan.hdr.n_loci, n_cells = (uint32_t)plan.hdr.n_cells;
    // CUDA start-up (driver + context, ~1 s on a B200 box) runs on a second thread while this one parses; the COO
    // is parsed into pinned memory when the context is already up, else into pageable memory (clx_load_coo takes both)
    clx_ctx* ctx = nullptr;
    int create_rc = CLX_OK;
    std::string create_err;
    std::thread cuda_init;
    if (!params.parse_only && !params.partition_only && params.gpus == 1)
        cuda_init = std::thread([&] {
            create_rc = clx_create(&ctx, params.device, params.has_vcf ? CLX_FLAG_KEEP_COO : 0u);
            if (create_rc != CLX_OK) create_err = clx_last_error(nullptr);
        });
    const size_t coo_bytes = (size_t)std::max<uint64_t>(nnz0, 1) * 4 * sizeof(uint32_t);
    uint32_t* coo = (uint32_t*)malloc(coo_bytes);
    if (!coo) die("out of host memory");
    CooSink sink{coo, coo + nnz0, coo + 2 * nnz0, coo + 3 * nnz0};
    parse_mtx_pair(alt, ref, plan, sink, params.threads);
    const double t_parse1 = now_s();
    if (params.parse_only) {
        printf("parse_only loci %u cells %u pairs %llu fnv_locus %llu fnv_cell %llu fnv_ref %llu fnv_alt %llu seconds %.6f threads %u\n", n_loci0,
               n_cells, (unsigned long long)nnz0, (unsigned long long)fnv1a(sink.locus, nnz0), (unsigned long long)fnv1a(sink.cell, nnz0),
               (unsigned long long)fnv1a(sink.ref, nnz0), (unsigned long long)fnv1a(sink.alt, nnz0), t_parse1 - t_parse0, params.threads);
        free(coo);
        return 0;
    }
    // init_cell_data indexes both vectors by cell id (load_data.rs:228-231): too few lines is a panic there
    if (barcodes.size() < n_cells || ground_truth.size() < n_cells)
        die("index out of bounds: the barcodes file has " + std::to_string(barcodes.size()) + " lines (" +
            std::to_string(ground_truth.size()) + " distinct) but the matrix has " + std::to_string(n_cells) + " cells");
    std::vector<VcfLocus> vcf_data;
    if (params.has_vcf) vcf_data = load_vcf_data(params);

    // ---- per-cell result arrays: plain memory on one GPU, a shared mapping the ranks fill side by side on several
    const size_t N = n_cells;
    const size_t cell_bytes = N * (7 * sizeof(double) + sizeof(uint32_t) + 3) + 64;
    unsigned char* cell_mem = params.gpus > 1 ? (unsigned char*)shared_alloc(cell_bytes) : (unsigned char*)calloc(cell_bytes, 1);
    if (!cell_mem) die("out of host memory");
    double* const ll = (double*)cell_mem;
    double *const ell = ll + N, *const norm = ell + N, *const post = norm + N, *const dpost = post + N, *const ll_maj = dpost + N,
                  *const ll_min = ll_maj + N;
    uint32_t* const nloci = (uint32_t*)(ll_min + N);
    uint8_t *const label = (uint8_t*)(nloci + N), *const anomaly = label + N, *const qual = anomaly + N;

    // ---- --gpus N: contiguous cell ranges with equal shares of the nonzeros, one process per GPU
    uint32_t c0 = 0, c1 = n_cells;  // this rank's cells
    if (params.gpus > 1) {
        if (n_cells < (uint32_t)params.gpus) die("--gpus: fewer cells than GPUs");
        std::vector<uint64_t> start(n_cells + 1, 0);
        for (uint64_t i = 0; i < nnz0; ++i) {
            if (sink.cell[i] >= n_cells) die("cell index out of range in the matrix");
            start[sink.cell[i] + 1]++;
        }
        for (uint32_t c = 0; c < n_cells; ++c) start[c + 1] += start[c];
        std::vector<uint32_t> bound(params.gpus + 1, n_cells);
        bound[0] = 0;
        for (int r = 1; r < params.gpus; ++r) {
            const uint64_t want = nnz0 / params.gpus * r;
            uint32_t b = (uint32_t)(std::lower_bound(start.begin(), start.end(), want) - start.begin());
            b = std::max(b, bound[r - 1] + 1);                                  // every rank gets a cell
            bound[r] = std::min(b, n_cells - (uint32_t)(params.gpus - r));
        }
        if (params.partition_only) {  // host self-check of the --gpus split (no CUDA): cell range and entries of every rank
            for (int r = 0; r < params.gpus; ++r)
                printf("partition rank %d cells %u %u entries %llu\n", r, bound[r], bound[r + 1],
                       (unsigned long long)(start[bound[r + 1]] - start[bound[r]]));
            free(coo);
            return 0;
        }
        g_team = (TeamShared*)shared_alloc(sizeof(TeamShared));
        new (g_team) TeamShared();
        g_world = params.gpus;
        fflush(stdout);
        fflush(stderr);
        for (int r = 1; r < params.gpus; ++r) {
            const pid_t pid = fork();  // no CUDA call has been made yet: every process initialises its own GPU
            if (pid < 0) die("fork failed");
            if (pid == 0) {
                g_rank = r;
                g_children.clear();
                prctl(PR_SET_PDEATHSIG, SIGKILL);
                break;
            }
            g_children.push_back(pid);
        }
        c0 = bound[g_rank];
        c1 = bound[g_rank + 1];
    }
    if (params.partition_only) {  // --gpus 1: one range
        printf("partition rank 0 cells 0 %u entries %llu\n", n_cells, (unsigned long long)nnz0);
        free(coo);
        return 0;
    }
    const bool root = g_rank == 0;
    const uint32_t n_local = c1 - c0;

    // ---- device: load_cell_data
    const double t_join0 = now_s();
    if (cuda_init.joinable()) cuda_init.join();
    else if (!params.parse_only) {
        create_rc = clx_create(&ctx, params.device + g_rank, params.has_vcf ? CLX_FLAG_KEEP_COO : 0u);
        if (create_rc != CLX_OK) create_err = clx_last_error(nullptr);
    }
    const double t_cuda_wait = now_s() - t_join0;
    if (create_rc != CLX_OK) die("clx_create failed: " + create_err);
    uint32_t* lcoo = nullptr;  // this rank's entries, file order kept, cell ids rank-local
    uint64_t nnz0_local = nnz0;
    if (g_world > 1) {
        if (root) {
            CLX(clx_nccl_unique_id(g_team->nccl_id));
            g_team->id_ready.store(1);
        }
        for (uint64_t spins = 0; !g_team->id_ready.load(); ++spins) {
            if (g_team->failed.load() || spins > 3000000) die("no NCCL id from rank 0");
            usleep(100);
        }
        CLX(clx_comm_init(ctx, g_rank, g_world, g_team->nccl_id));
        nnz0_local = 0;
        for (uint64_t i = 0; i < nnz0; ++i) nnz0_local += sink.cell[i] >= c0 && sink.cell[i] < c1;
        lcoo = (uint32_t*)malloc((size_t)std::max<uint64_t>(nnz0_local, 1) * 4 * sizeof(uint32_t));
        if (!lcoo) die("out of host memory");
        uint64_t k = 0;
        for (uint64_t i = 0; i < nnz0; ++i)
            if (sink.cell[i] >= c0 && sink.cell[i] < c1) {
                lcoo[k] = sink.locus[i];
                lcoo[nnz0_local + k] = sink.cell[i] - c0;
                lcoo[2 * nnz0_local + k] = sink.ref[i];
                lcoo[3 * nnz0_local + k] = sink.alt[i];
                ++k;
            }
    }
    const double t_load0 = now_s();
    uint32_t n_loci = 0;
    uint64_t nnz = 0;
    if (params.min_ref > 0xFFFFFFFFull || params.min_alt > 0xFFFFFFFFull) die("min_ref / min_alt out of range");
    if (g_world > 1)
        CLX(clx_load_coo(ctx, nnz0_local, n_loci0, n_local, lcoo, lcoo + nnz0_local, lcoo + 2 * nnz0_local, lcoo + 3 * nnz0_local,
                         (uint32_t)params.min_ref, (uint32_t)params.min_alt, &n_loci, &nnz));
    else
        CLX(clx_load_coo(ctx, nnz0, n_loci0, n_cells, sink.locus, sink.cell, sink.ref, sink.alt, (uint32_t)params.min_ref,
                         (uint32_t)params.min_alt, &n_loci, &nnz));
    free(lcoo);
    std::vector<uint32_t> locus_ids(n_loci);
    CLX(clx_get_locus_ids(ctx, locus_ids.data()));
    free(coo);
    coo = nullptr;
    const double t_load1 = now_s();

    // ---- cellector(): the EM loop (main.rs:36-46)
    std::vector<double> l_min(n_loci), l_maj(n_loci);
    std::vector<uint32_t> c_min(n_loci), c_maj(n_loci);
    std::vector<uint64_t> r_min(n_loci), a_min(n_loci), r_maj(n_loci), a_maj(n_loci);
    std::vector<double> pc_min(n_loci), pc_maj(n_loci);
    std::vector<uint32_t> ordering(n_loci);
    std::vector<double> em_device_ms;
    double t_em = 0.0, t_write = 0.0;
    uint64_t iteration = 0;
    for (;;) {
        const double t0 = now_s();
        clx_iter_stats st;
        CLX(clx_em_iteration(ctx, params.interquartile_range_multiple, &st));
        em_device_ms.push_back(st.device_ms);
        const double t1 = now_s();
        t_em += t1 - t0;
        // main.rs:338-339
        if (root) {
            printf("detected %llu new anomylous cells and rescued %llu cells to the majority in iteration %llu\n",
                   (unsigned long long)st.n_new_excluded, (unsigned long long)st.n_rescued, (unsigned long long)(iteration + 1));
            printf("median normalized log likelihood %s with interquartile range %s, threshold %s\n", f64_str(st.median).c_str(),
                   f64_str(st.iqr).c_str(), f64_str(st.threshold).c_str());
        }
        const bool files = !params.no_iteration_files;
        if (files) {  // every rank puts its cells' vectors side by side; rank 0 writes them below
            CLX(clx_get_cell_vectors(ctx, ll + c0, ell + c0, nloci + c0, norm + c0, nullptr));
            team_barrier();
        }
        if (root && (files || st.n_loci_filtered > 0)) {  // the per-locus vectors are the same on every rank
            clx_locus_arrays la;
            memset(&la, 0, sizeof(la));
            la.ll_minority = l_min.data();
            la.ll_majority = l_maj.data();
            la.cells_minority = c_min.data();
            la.cells_majority = c_maj.data();
            la.ref_minority = r_min.data();
            la.alt_minority = a_min.data();
            la.ref_majority = r_maj.data();
            la.alt_majority = a_maj.data();
            CLX(clx_get_locus_vectors(ctx, &la));
            // locus_filter_and_output_locus_data (main.rs:422-498)
            std::vector<double> for_threshold;
            for (uint32_t i = 0; i < n_loci; ++i) {
                if (c_min[i] != 0) {
                    pc_min[i] = l_min[i] / (double)c_min[i];
                    for_threshold.push_back(pc_min[i]);
                } else pc_min[i] = 0.0;
                pc_maj[i] = c_maj[i] != 0 ? l_maj[i] / (double)c_maj[i] : 0.0;
                if (pc_min[i] != pc_min[i]) die("called `Option::unwrap()` on a `None` value (NaN in argsort, main.rs:504)");
            }
            if (st.n_loci_filtered > 0) {
                const double median = median_of(for_threshold);
                for (uint32_t i = 0; i < n_loci; ++i)
                    if (pc_min[i] < -80.0)
                        printf("filtering locus %u locus index %u because it was contributing %s vs median %s per cell to log likelihood of minority cells\n",
                               locus_ids[i], i, f64_str(pc_min[i]).c_str(), f64_str(median).c_str());
            }
            if (files) {
                std::iota(ordering.begin(), ordering.end(), 0u);
                std::stable_sort(ordering.begin(), ordering.end(), [&](uint32_t a, uint32_t b) { return pc_min[a] < pc_min[b]; });
                FileWriter w(params.output_directory + "/iteration_" + std::to_string(iteration) + "_locus_contribution.tsv");
                w.buf +=
                    "locus_id\tchrom\tpos\tlog_likelihood_minority\tlog_likelihood_majority\texpected_loglike_minority\texpected_loglike_"
                    "majority\tminority_cellcount\tmajority_cellcount\tlog_likelihood_minority_per_cell\tlog_likelihood_majority_per_cell\t"
                    "minority_alt\tminority_ref\tmajority_alt\tmajority_ref\tminority_af\tmajority_af\n";
                for (const uint32_t i : ordering) {
                    const uint32_t locus_id = locus_ids[i];
                    double min_af = 0.0, maj_af = 0.0;
                    if (a_min[i] + r_min[i] > 0) min_af = (double)a_min[i] / (double)(a_min[i] + r_min[i]);
                    if (a_maj[i] + r_maj[i] > 0) maj_af = (double)a_maj[i] / (double)(a_maj[i] + r_maj[i]);
                    std::string& b = w.buf;
                    append_u64(b, locus_id);
                    b += '\t';
                    if (params.has_vcf) {
                        if (locus_id >= vcf_data.size()) die("index out of bounds: locus " + std::to_string(locus_id) + " has no vcf record");
                        b += vcf_data[locus_id].chrom;
                        b += '\t';
                        b += vcf_data[locus_id].pos;
                    } else b += "na\tna";
                    b += '\t';
                    append_f64(b, l_min[i]); b += '\t';
                    append_f64(b, l_maj[i]); b += '\t';
                    append_f64(b, l_min[i]); b += '\t';  // "expected" columns repeat the actual ones (main.rs:394)
                    append_f64(b, l_maj[i]); b += '\t';
                    append_u64(b, c_min[i]); b += '\t';
                    append_u64(b, c_maj[i]); b += '\t';
                    append_f64(b, pc_min[i]); b += '\t';
                    append_f64(b, pc_maj[i]); b += '\t';
                    append_u64(b, a_min[i]); b += '\t';
                    append_u64(b, r_min[i]); b += '\t';
                    append_u64(b, a_maj[i]); b += '\t';
                    append_u64(b, r_maj[i]); b += '\t';
                    append_f64(b, min_af); b += '\t';
                    append_f64(b, maj_af); b += '\n';
                    w.maybe_flush();
                }
            }
        }
        if (files && root) {  // output_iteration_tsv (main.rs:349-366)
            FileWriter w(params.output_directory + "/iteration_" + std::to_string(iteration) + ".tsv");
            w.buf += "cell_id\tbarcode\tassignment\tlog_likelihood\texpected_log_likelihood\tnum_loci_used\n";
            for (uint32_t c = 0; c < n_cells; ++c) {
                std::string& b = w.buf;
                append_u64(b, c); b += '\t';
                b += barcodes[c]; b += '\t';
                b += ground_truth[c]; b += '\t';
                append_f64(b, ll[c]); b += '\t';
                append_f64(b, ell[c]); b += '\t';
                append_u64(b, nloci[c]); b += '\n';  // an f64 holding an integer prints without ".0"
                w.maybe_flush();
            }
            FileWriter wt(params.output_directory + "/iteration_" + std::to_string(iteration) + "_threshold.tsv");
            append_f64(wt.buf, st.threshold);
        }
        t_write += now_s() - t1;
        ++iteration;
        if (!st.any_change) break;
        if (params.max_iterations && iteration >= params.max_iterations) break;
    }
    if (params.no_iteration_files) CLX(clx_get_cell_vectors(ctx, nullptr, nullptr, nloci + c0, norm + c0, nullptr));

    // ---- calculate_posteriors (main.rs:228-280) + assignment rule
    const double t_post0 = now_s();
    if (params.min_loci_used > 0xFFFFFFFFull) die("min_loci_for_assignment out of range");
    CLX(clx_posteriors(ctx, params.posterior_threshold, (uint32_t)params.min_loci_used, post + c0, dpost + c0, ll_maj + c0,
                       ll_min + c0, label + c0, anomaly + c0, qual + c0));
    const double t_post1 = now_s();

    // ---- output_final_vcf (main.rs:52-131)
    std::vector<uint64_t> mn, mj;
    if (params.has_vcf) {  // the counts are summed over the ranks inside the call: every rank makes it
        mn.resize(2 * (size_t)n_loci0 + 2);
        mj.resize(2 * (size_t)n_loci0 + 2);
        CLX(clx_final_locus_counts(ctx, mn.data(), mj.data()));
    }
    team_barrier();  // every rank's posterior arrays are in place
    if (!root) {     // workers are done: rank 0 writes the remaining files
        clx_destroy(ctx);
        fflush(nullptr);
        _exit(0);
    }
    if (params.has_vcf) {
        FileWriter w(params.output_directory + "/cellector.vcf");
        size_t record = 0;
        const double ambient = 0.03, gt_threshold = 0.99;
        for (const std::string& line : read_lines(params.vcf)) {
            std::string& b = w.buf;
            if (line.rfind("##", 0) == 0) {
                b += line;
                b += '\n';
            } else if (line.rfind("#CHROM", 0) == 0) {
                b += line;
                b += "\tmajority\tminority\n";
            } else {
                if (record >= n_loci0) die("index out of bounds: the vcf has more records than the matrix has loci");
                const uint64_t min_ref = mn[2 * record], min_alt = mn[2 * record + 1], maj_ref = mj[2 * record], maj_alt = mj[2 * record + 1];
                const uint64_t total_alt = min_alt + maj_alt, total_ref = min_ref + maj_ref;
                double soup = 0.5;
                if (total_alt + total_ref > 0) soup = (double)total_alt / (double)(total_alt + total_ref);
                const double p_hom_alt = (1.0 - ambient) * 0.99 + ambient * soup;
                const double p_het = (1.0 - ambient) * 0.5 + ambient * soup;
                const double p_hom_ref = (1.0 - ambient) * 0.01 + ambient * soup;
                auto call = [&](uint64_t a, uint64_t r, const char** gt, double* max_post) {
                    const double l_aa = binomial_pmf(p_hom_alt, a + r, a), l_het = binomial_pmf(p_het, a + r, a),
                                 l_rr = binomial_pmf(p_hom_ref, a + r, a);
                    const double denom = 1.0 / 3.0 * l_aa + 1.0 / 3.0 * l_het + 1.0 / 3.0 * l_rr;
                    const double p_aa = l_aa * 1.0 / 3.0 / denom, p_h = l_het * 1.0 / 3.0 / denom, p_rr = l_rr * 1.0 / 3.0 / denom;
                    *max_post = fmax(fmax(p_aa, p_h), p_rr);  // f64::max ignores a NaN operand like fmax
                    *gt = "./.";
                    if (p_aa > gt_threshold) *gt = "1/1";
                    else if (p_h > gt_threshold) *gt = "0/1";
                    else if (p_rr > gt_threshold) *gt = "0/0";
                };
                const char *gt_maj, *gt_min;
                double mp_maj, mp_min;
                call(min_alt, min_ref, &gt_min, &mp_min);
                call(maj_alt, maj_ref, &gt_maj, &mp_maj);
                b += line;
                b += "\tGT:GP:AO:RO\t";
                b += gt_maj; b += ':'; append_f64(b, mp_maj); b += ':'; append_u64(b, maj_alt); b += ':'; append_u64(b, maj_ref);
                b += '\t';
                b += gt_min; b += ':'; append_f64(b, mp_min); b += ':'; append_u64(b, min_alt); b += ':'; append_u64(b, min_ref);
                b += '\n';
                ++record;
            }
            w.maybe_flush();
        }
    }

    // ---- output_final_assignments (main.rs:133-174)
    const double t_out0 = now_s();
    {
        static const char* label_str[4] = {"0", "1", "doublet", "unassigned"};
        FileWriter w(params.output_directory + "/cellector_assignments.tsv");
        w.buf +=
            "barcode\tposterior_assignment\tanomally_assignment\tlog_likelihood_loci_normalized\tloci_used\tposterior_assign_qual\t"
            "majority_log_likelihood\tminority_log_likelihood\tground_truth_assignment\n";
        std::map<std::string, std::map<std::string, uint64_t>> assignment_gt_counts;
        std::map<std::string, uint64_t> gt_counts;
        for (uint32_t c = 0; c < n_cells; ++c) {
            const char* pa = label_str[label[c] & 3];
            assignment_gt_counts[pa][ground_truth[c]] += 1;
            gt_counts[ground_truth[c]] += 1;
            std::string& b = w.buf;
            b += barcodes[c]; b += '\t';
            b += pa; b += '\t';
            b += anomaly[c] ? "1" : "0"; b += '\t';
            append_f64(b, norm[c]); b += '\t';
            append_u64(b, nloci[c]); b += '\t';
            append_u64(b, qual[c]); b += '\t';
            append_f64(b, ll_maj[c]); b += '\t';
            append_f64(b, ll_min[c]); b += '\t';
            b += ground_truth[c]; b += '\n';
            w.maybe_flush();
        }
        pretty_print(assignment_gt_counts, gt_counts);
    }
    const double t_end = now_s();
    t_write += t_end - t_out0;
    uint64_t launches = 0;
    clx_launch_count(ctx, &launches);
    if (!params.timing_json.empty()) {
        FileWriter w(params.timing_json);
        std::string& b = w.buf;
        b += "{\"parse_s\": " + f64_str(t_parse1 - t_parse0) + ", \"cuda_init_wait_s\": " + f64_str(t_cuda_wait) + ", \"load_s\": " + f64_str(t_load1 - t_load0) + ", \"em_s\": " + f64_str(t_em) +
             ", \"posterior_s\": " + f64_str(t_post1 - t_post0) + ", \"write_s\": " + f64_str(t_write) + ", \"total_s\": " + f64_str(t_end - t_start) +
             ", \"iterations\": " + std::to_string(iteration) + ", \"nnz0\": " + std::to_string(nnz0) + ", \"nnz_used\": " + std::to_string(nnz) +
             ", \"n_loci_used\": " + std::to_string(n_loci) + ", \"n_cells\": " + std::to_string(n_cells) + ", \"gpus\": " +
             std::to_string(g_world) + ", \"threads\": " +
             std::to_string(params.threads) + ", \"gpu_launches\": " + std::to_string(launches) + ", \"em_device_ms\": [";
        for (size_t i = 0; i < em_device_ms.size(); ++i) b += (i ? ", " : "") + f64_str(em_device_ms[i]);
        b += "]}\n";
    }
    clx_destroy(ctx);
    for (const pid_t pid : g_children) {
        int st = 0;
        waitpid(pid, &st, 0);
        if (!(WIFEXITED(st) && WEXITSTATUS(st) == 0)) die("a worker process ended abnormally");
    }
    return 0;
}
