// souporcell — host CLI over libcellector_cuda.so for souporcell's clustering step: the reference's command line and
// stdout (souporcell/souporcell/src/params.yml, souporcell/souporcell/src/main.rs:35-106) with the k-cluster EM on the
// GPU behind clx_soup_restarts (include/cellector_cuda.h).  cellector_pipeline.py:233-240 runs it as
//   souporcell -a alt.mtx -r ref.mtx --barcodes barcodes.tsv -k 2 [--restarts N] [-t threads] [--seed S] > souporcell.out
// and feeds the output (barcode, best cluster, one log-probability per cluster) to troublet.
//
//   souporcell -a alt.mtx -r ref.mtx -b barcodes.tsv -k K [--min_alt 4] [--min_ref 4] [--restarts 100] [-t 1] [--seed 4]
//              [--initialization_strategy random_uniform]
//              [-g known_genotypes.vcf(.gz) --known_genotypes_sample_names S0 S1 ...]
//   additive flags: --device N, --parse_threads N, --rng_selftest
//   refused (not built): --known_cell_assignments, initialization
//   strategies other than random_uniform (the reference itself asserts on kmeans++, middle_variance and
//   known_cell_assignments: main.rs:341-349, 389-392; random_cell_assignment needs rand's gen_range).
//
// Differences a user can see: the per-iteration `binomial ...` and per-restart `thread ... done` lines on stderr are
// not printed (the loop state lives on the device); the numbers come from f64 sums and are printed as f32 like the
// reference's, so last digits can differ from the Rust binary's f32 arithmetic.
#include <unistd.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "../../../include/cellector_cuda.h"
#include "mtx_parse.h"
#include "rust_fmt.h"
#include "soup_rng.h"

using namespace clxhost;

namespace {

[[noreturn]] void die(const std::string& msg, int code = 101) {  // 101 = exit status of a Rust panic
    fprintf(stderr, "souporcell: %s\n", msg.c_str());
    exit(code);
}

// Rust `{}` for f32: shortest digits that round-trip as f32, positional
void append_f32(std::string& out, float v) {
    if (std::isnan(v)) {
        out += "NaN";
        return;
    }
    if (std::isinf(v)) {
        out += v < 0 ? "-inf" : "inf";
        return;
    }
    char buf[64];
    const auto r = std::to_chars(buf, buf + sizeof(buf), v, std::chars_format::scientific);
    const char* p = buf;
    if (*p == '-') {
        out += '-';
        ++p;
    }
    char digits[16];
    int nd = 0;
    const char* q = p;
    for (; q < r.ptr && *q != 'e'; ++q)
        if (*q != '.') digits[nd++] = *q;
    int exp10 = 0;
    std::from_chars(q + 1 + (q[1] == '+' ? 1 : 0), r.ptr, exp10);
    while (nd > 1 && digits[nd - 1] == '0') --nd;
    const int point = exp10 + 1;
    if (point <= 0) {
        out += "0.";
        out.append((size_t)(-point), '0');
        out.append(digits, (size_t)nd);
    } else if (point >= nd) {
        out.append(digits, (size_t)nd);
        out.append((size_t)(point - nd), '0');
    } else {
        out.append(digits, (size_t)point);
        out += '.';
        out.append(digits + point, (size_t)(nd - point));
    }
}

struct Params {  // main.rs:494-510
    std::string ref_mtx, alt_mtx, barcodes;
    uint64_t num_clusters = 0, min_alt = 4, min_ref = 4, min_alt_umis = 0, min_ref_umis = 0, restarts = 100, threads = 1;
    uint8_t seed = 4;
    bool has_known_genotypes = false;
    std::string known_genotypes;
    std::vector<std::string> known_genotypes_sample_names;
    int device = 0;
    unsigned parse_threads = 0;
};

uint64_t parse_u64(const std::string& name, const std::string& v) {
    const char* p = v.data();
    uint64_t out = 0;
    if (!next_usize(p, v.data() + v.size(), &out) || p != v.data() + v.size()) die("invalid value '" + v + "' for --" + name);
    return out;
}

Params load_params(int argc, char** argv) {
    static const std::map<std::string, std::string> shorts = {{"-r", "ref_matrix"}, {"-a", "alt_matrix"}, {"-b", "barcodes"},
                                                             {"-k", "num_clusters"}, {"-g", "known_genotypes"}, {"-t", "threads"}};
    std::map<std::string, std::string> kv;
    std::vector<std::string> sample_names;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i], name, val;
        bool have_val = false;
        if (a.rfind("--", 0) == 0) {
            const size_t eq = a.find('=');
            name = a.substr(2, eq == std::string::npos ? std::string::npos : eq - 2);
            if (eq != std::string::npos) {
                val = a.substr(eq + 1);
                have_val = true;
            }
        } else if (a.size() >= 2 && a[0] == '-' && shorts.count(a.substr(0, 2))) {
            name = shorts.at(a.substr(0, 2));
            if (a.size() > 2) {
                val = a.substr(a[2] == '=' ? 3 : 2);
                have_val = true;
            }
        } else {
            die("error: Found argument '" + a + "' which wasn't expected, or isn't valid in this context", 1);
        }
        if (!have_val) {
            if (i + 1 >= argc) die("error: The argument '--" + name + "' requires a value but none was supplied", 1);
            val = argv[++i];
        }
        if (name == "known_genotypes_sample_names") {  // `multiple: true` (params.yml:53-57): every value up to the next flag
            sample_names.push_back(val);
            while (i + 1 < argc && argv[i + 1][0] != '-') sample_names.push_back(argv[++i]);
        }
        kv[name] = val;
    }
    static const char* known[] = {"ref_matrix", "alt_matrix", "barcodes", "num_clusters", "min_alt", "min_ref", "restarts",
                                  "known_genotypes", "known_genotypes_sample_names", "known_cell_assignments",
                                  "initialization_strategy", "threads", "seed", "min_alt_umis", "min_ref_umis", "device",
                                  "parse_threads"};
    for (const auto& e : kv) {
        bool ok = false;
        for (const char* k : known) ok = ok || e.first == k;
        if (!ok) die("error: Found argument '--" + e.first + "' which wasn't expected, or isn't valid in this context", 1);
    }
    for (const char* req : {"ref_matrix", "alt_matrix", "barcodes", "num_clusters"})
        if (!kv.count(req)) die(std::string("error: The following required arguments were not provided:\n    --") + req, 1);
    Params p;
    p.ref_mtx = kv["ref_matrix"];
    p.alt_mtx = kv["alt_matrix"];
    p.barcodes = kv["barcodes"];
    p.num_clusters = parse_u64("num_clusters", kv["num_clusters"]);
    if (kv.count("min_alt")) p.min_alt = parse_u64("min_alt", kv["min_alt"]);
    if (kv.count("min_ref")) p.min_ref = parse_u64("min_ref", kv["min_ref"]);
    if (kv.count("restarts")) p.restarts = parse_u64("restarts", kv["restarts"]);
    if (kv.count("threads")) p.threads = parse_u64("threads", kv["threads"]);
    if (kv.count("seed")) {
        const uint64_t s = parse_u64("seed", kv["seed"]);
        if (s > 255) die("called `Result::unwrap()` on an `Err` value: ParseIntError { kind: PosOverflow } (seed is a u8, main.rs:584)");
        p.seed = (uint8_t)s;
    }
    if (kv.count("min_alt_umis")) p.min_alt_umis = parse_u64("min_alt_umis", kv["min_alt_umis"]);
    if (kv.count("min_ref_umis")) p.min_ref_umis = parse_u64("min_ref_umis", kv["min_ref_umis"]);
    if (kv.count("device")) p.device = (int)parse_u64("device", kv["device"]);
    if (kv.count("parse_threads")) p.parse_threads = (unsigned)parse_u64("parse_threads", kv["parse_threads"]);
    if (p.parse_threads == 0) p.parse_threads = std::max(1u, std::thread::hardware_concurrency());
    if (kv.count("known_genotypes")) {
        p.has_known_genotypes = true;
        p.known_genotypes = kv["known_genotypes"];
    }
    p.known_genotypes_sample_names = sample_names;
    if (kv.count("known_cell_assignments")) die("known cell assignments not yet implemented");  // the reference's assert, main.rs:342
    if (kv.count("initialization_strategy") && kv["initialization_strategy"] != "random_uniform") {
        const std::string& s = kv["initialization_strategy"];
        if (s == "kmeans++") die("kmeans++ not yet implemented");                  // main.rs:347
        if (s == "middle_variance") die("middle variance not yet implemented");    // main.rs:390
        if (s == "random_cell_assignment") die("initialization_strategy random_cell_assignment is not built here (main.rs:366-387)");
        die("initialization strategy must be one of kmeans++, random_uniform, random_cell_assignment, middle_variance");
    }
    if ((p.min_alt_umis != 0 || p.min_ref_umis != 0) && p.min_alt == 0 && p.min_ref == 0)
        die("--min_alt_umis / --min_ref_umis need --min_alt or --min_ref of at least 1 in this build (a locus that fails the UMI "
            "minima is dropped on the host before clx_load_coo, whose own filter then has to reject the emptied locus)");
    if (p.num_clusters < 1 || p.num_clusters > 8) die("--num_clusters must be between 1 and 8 in this build");
    if (p.threads < 1) die("attempt to divide by zero (threads = 0)");
    return p;
}

// init_cluster_centers_known_genotypes (main.rs:309-338): 0.5 everywhere, then for every VCF record whose 0-based record
// number is a used locus, cluster s takes (hap0 + hap1) / 2 of sample s (alleles capped at 1, clamped to [0.01, 0.99]);
// a GT starting with '.' leaves 0.5.  Characters 0 and 2 of the GT string, like the reference.
std::vector<float> known_genotype_centers(const Params& p, uint32_t k, uint32_t n_loci, const std::vector<uint32_t>& locus_ids) {
    if (p.known_genotypes_sample_names.empty()) die("currently requiring known_genotypes_sample_names if known_genotypes set");
    if (p.known_genotypes_sample_names.size() > k)
        die("index out of bounds: the len is " + std::to_string(k) + " but the index is " + std::to_string(k) +
            " (more sample names than clusters, main.rs:331)");
    std::vector<float> centers((size_t)k * n_loci, 0.5f);
    std::vector<int64_t> index_of;  // record number -> used-locus index
    for (uint32_t i = 0; i < n_loci; ++i) {
        if (locus_ids[i] >= index_of.size()) index_of.resize((size_t)locus_ids[i] + 1, -1);
        index_of[locus_ids[i]] = i;
    }
    FileBytes f(p.known_genotypes);
    std::vector<int> sample_col;  // column of every requested sample
    size_t record = 0;
    for_each_line(f.data(), f.size(), [&](const char* s, size_t n) {
        if (n >= 2 && s[0] == '#' && s[1] == '#') return;
        std::vector<std::pair<const char*, size_t>> col;
        for (size_t b = 0; b <= n;) {
            size_t e = b;
            while (e < n && s[e] != '\t') ++e;
            col.emplace_back(s + b, e - b);
            b = e + 1;
        }
        if (n >= 1 && s[0] == '#') {  // #CHROM line: sample names from column 9 on
            for (const std::string& name : p.known_genotypes_sample_names) {
                int at = -1;
                for (size_t c = 9; c < col.size(); ++c)
                    if (std::string(col[c].first, col[c].second) == name) at = (int)c;
                if (at < 0) die("sample " + name + " is not in the known_genotypes vcf");
                sample_col.push_back(at);
            }
            return;
        }
        if (n == 0) return;
        const size_t this_record = record++;
        if (this_record >= index_of.size() || index_of[this_record] < 0) return;
        if (sample_col.empty()) die("the known_genotypes vcf has no #CHROM header line");
        if (col.size() < 10) die("vcf record without sample columns");
        int gt_field = -1, fi = 0;  // position of GT in the FORMAT column
        for (size_t b = 0; b <= col[8].second; ++fi) {
            size_t e = b;
            while (e < col[8].second && col[8].first[e] != ':') ++e;
            if (e - b == 2 && col[8].first[b] == 'G' && col[8].first[b + 1] == 'T') gt_field = fi;
            b = e + 1;
        }
        if (gt_field < 0) die("vcf record without a GT field");
        for (size_t si = 0; si < sample_col.size(); ++si) {
            if ((size_t)sample_col[si] >= col.size()) die("vcf record with too few sample columns");
            const char* v = col[sample_col[si]].first;
            const size_t vn = col[sample_col[si]].second;
            size_t b = 0;
            for (int skip = 0; skip < gt_field; ++skip) {
                while (b < vn && v[b] != ':') ++b;
                ++b;
            }
            size_t e = b;
            while (e < vn && v[e] != ':') ++e;
            if (b >= vn || e == b) die("empty GT");
            if (v[b] == '.') continue;  // main.rs:324
            if (e - b < 3) die("called `Option::unwrap()` on a `None` value (GT shorter than three characters, main.rs:326)");
            if (v[b] < '0' || v[b] > '9' || v[b + 2] < '0' || v[b + 2] > '9') die("called `Result::unwrap()` on an `Err` value: ParseIntError (GT)");
            const unsigned hap0 = std::min(1u, (unsigned)(v[b] - '0')), hap1 = std::min(1u, (unsigned)(v[b + 2] - '0'));
            centers[si * (size_t)n_loci + (size_t)index_of[this_record]] = std::fmax(std::fmin((float)(hap0 + hap1) / 2.0f, 0.99f), 0.01f);
        }
    });
    return centers;
}

#define CLX(call)                                                                                  \
    do {                                                                                           \
        const int rc__ = (call);                                                                   \
        if (rc__ != CLX_OK) die(std::string(#call) + " failed: " + clx_last_error(ctx));           \
    } while (0)

// main.rs:89-105: the best loss on stderr; zip(barcodes, log probabilities): barcode, first best cluster, every log probability
void print_results(const std::vector<std::string>& barcodes, const std::vector<double>& log_probs, uint32_t n_cells, uint32_t k,
                   double best_loss) {
    {
        std::string s = "best total log probability = ";
        append_f32(s, (float)best_loss);  // main.rs:89
        fprintf(stderr, "%s\n", s.c_str());
    }
    // main.rs:91-105: zip(barcodes, log probabilities): barcode, first best cluster, every log probability
    std::string out;
    out.reserve(1 << 20);
    const size_t n_out = std::min<size_t>(barcodes.size(), n_cells);
    for (size_t c = 0; c < n_out; ++c) {
        const double* lp = &log_probs[c * k];
        uint32_t bi = 0;
        float blp = -INFINITY;
        for (uint32_t q = 0; q < k; ++q)
            if ((float)lp[q] > blp) {
                bi = q;
                blp = (float)lp[q];
            }
        out += barcodes[c];
        out += '\t';
        append_u64(out, bi);
        out += '\t';
        for (uint32_t q = 0; q < k; ++q) {
            append_f32(out, (float)lp[q]);
            if (q + 1 < k) out += '\t';
        }
        out += '\n';
        if (out.size() > (1 << 20) - 4096) {
            fwrite(out.data(), 1, out.size(), stdout);
            out.clear();
        }
    }
    fwrite(out.data(), 1, out.size(), stdout);
}

}  // namespace

int main(int argc, char** argv) {
    if (argc >= 2 && std::string(argv[1]) == "--rng_selftest") {
        // ChaCha20 block of RFC 7539 section 2.3.2: key 00..1f, counter 1, nonce 00:00:00:09:00:00:00:4a:00:00:00:00
        uint32_t st[16] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u}, out[16];
        for (int i = 0; i < 8; ++i)
            st[4 + i] = (uint32_t)(4 * i) | (uint32_t)(4 * i + 1) << 8 | (uint32_t)(4 * i + 2) << 16 | (uint32_t)(4 * i + 3) << 24;
        st[12] = 1;
        st[13] = 0x09000000u;
        st[14] = 0x4a000000u;
        st[15] = 0;
        chacha20_block(st, out);
        for (int i = 0; i < 16; ++i) printf("%08x%c", out[i], i == 15 ? '\n' : ' ');
        // the first draws of StdRng::from_seed([4; 32]) as this file maps them (u8, then f32)
        uint8_t seed[32];
        memset(seed, 4, sizeof(seed));
        StdRng07 rng = StdRng07::from_seed(seed);
        const unsigned b0 = rng.gen_u8(), b1 = rng.gen_u8(), b2 = rng.gen_u8();  // in this order (printf's is unspecified)
        printf("u8 %u %u %u f32", b0, b1, b2);
        std::string s;
        for (int i = 0; i < 3; ++i) {
            s += ' ';
            append_f32(s, rng.gen_f32());
        }
        printf("%s\n", s.c_str());
        s.clear();
        for (const float v : {0.1f, -2181.0225f, 5.0f, 1e-7f, 3.4028235e38f, -0.0f}) {
            append_f32(s, v);
            s += ' ';
        }
        printf("f32 %s\n", s.c_str());
        return 0;
    }
    const Params params = load_params(argc, argv);
    // load_barcodes (main.rs:478-489)
    std::vector<std::string> barcodes;
    {
        FileBytes f(params.barcodes);
        for_each_line(f.data(), f.size(), [&](const char* s, size_t n) { barcodes.emplace_back(s, n); });
    }
    // load_cell_data (main.rs:390-468): one parse of the pair, the rest on the device
    clx_ctx* ctx = nullptr;
    int create_rc = CLX_OK;
    std::string create_err;
    std::thread cuda_init([&] {
        create_rc = clx_create(&ctx, params.device, 0u);
        if (create_rc != CLX_OK) create_err = clx_last_error(nullptr);
    });
    FileBytes alt(params.alt_mtx), ref(params.ref_mtx);
    const MtxPairPlan plan = plan_mtx_pair(alt, ref, params.parse_threads);
    const uint64_t nnz0 = plan.n_pairs;
    if (plan.hdr.n_loci > 0xFFFFFFFEull || plan.hdr.n_cells > 0xFFFFFFFEull) die("matrix dimensions exceed 2^32 - 2");
    const uint32_t n_loci0 = (uint32_t)plan.hdr.n_loci, n_cells = (uint32_t)plan.hdr.n_cells;
    uint32_t* coo = (uint32_t*)malloc((size_t)std::max<uint64_t>(nnz0, 1) * 4 * sizeof(uint32_t));
    if (!coo) die("out of host memory");
    CooSink sink{coo, coo + nnz0, coo + 2 * nnz0, coo + 3 * nnz0};
    parse_mtx_pair(alt, ref, plan, sink, params.parse_threads);
    // the UMI minima of souporcell's locus filter (main.rs:440: summed ref / alt counts of the locus over all its lines):
    // clx_load_coo filters on cell counts only, so the lines of a locus that fails them are dropped here; the emptied
    // locus then fails min_ref / min_alt there, and the dense renumbering of the others is the same
    uint64_t nnz_load = nnz0;
    if (params.min_ref_umis != 0 || params.min_alt_umis != 0) {
        std::vector<uint64_t> umis(2 * (size_t)n_loci0, 0);
        for (uint64_t i = 0; i < nnz0; ++i) {
            if (sink.locus[i] >= n_loci0) die("assertion failed: locus < total_loci");  // main.rs:414
            umis[2 * (size_t)sink.locus[i]] += sink.ref[i];
            umis[2 * (size_t)sink.locus[i] + 1] += sink.alt[i];
        }
        uint64_t w = 0;
        for (uint64_t i = 0; i < nnz0; ++i) {
            const size_t l = sink.locus[i];
            if (umis[2 * l] >= params.min_ref_umis && umis[2 * l + 1] >= params.min_alt_umis) {
                sink.locus[w] = sink.locus[i];
                sink.cell[w] = sink.cell[i];
                sink.ref[w] = sink.ref[i];
                sink.alt[w] = sink.alt[i];
                ++w;
            }
        }
        nnz_load = w;
    }
    cuda_init.join();
    if (create_rc != CLX_OK) die("clx_create failed: " + create_err);
    uint32_t n_loci = 0;
    uint64_t nnz = 0;
    CLX(clx_load_coo(ctx, nnz_load, n_loci0, n_cells, sink.locus, sink.cell, sink.ref, sink.alt, (uint32_t)params.min_ref,
                     (uint32_t)params.min_alt, &n_loci, &nnz));
    free(coo);
    fprintf(stderr, "total loci used %u\n", n_loci);  // main.rs:465

    // souporcell_main (main.rs:62-88): one generator per thread seeded from the main one, solves_per_thread runs each;
    // a thread keeps its first best run, the threads are compared in order: the first maximum in thread-major order
    const uint32_t k = (uint32_t)params.num_clusters;
    if (params.has_known_genotypes) {
        // every restart of the reference starts from the same VCF-derived centers and ends in the same place (main.rs:303-305:
        // no draw is made); the first one is kept by the strict > of main.rs:73 — one run here
        std::vector<uint32_t> locus_ids(n_loci);
        CLX(clx_get_locus_ids(ctx, locus_ids.data()));
        const std::vector<float> centers = known_genotype_centers(params, k, n_loci, locus_ids);
        std::vector<double> log_probs((size_t)n_cells * k);
        double loss = 0.0;
        uint32_t iterations = 0;
        CLX(clx_soup_em(ctx, k, centers.data(), log_probs.data(), nullptr, &loss, &iterations));
        print_results(barcodes, log_probs, n_cells, k, loss);
        clx_destroy(ctx);
        return 0;
    }
    const uint64_t solves_per_thread = (params.restarts + params.threads - 1) / params.threads;  // (restarts / threads).ceil()
    const uint64_t runs = solves_per_thread * params.threads;
    if (runs == 0) die("no restarts");
    if (runs > 0xFFFFFFFFull) die("too many restarts");
    uint8_t seed[32];
    memset(seed, params.seed, sizeof(seed));
    StdRng07 main_rng = StdRng07::from_seed(seed);
    std::vector<float> centers((size_t)runs * k * n_loci);
    size_t w = 0;
    for (uint64_t t = 0; t < params.threads; ++t) {
        uint8_t ts[32];
        main_rng.new_seed(ts);
        StdRng07 rng = StdRng07::from_seed(ts);
        for (uint64_t it = 0; it < solves_per_thread; ++it)
            for (size_t i = 0; i < (size_t)k * n_loci; ++i)  // init_cluster_centers_uniform (main.rs:354-364)
                centers[w++] = std::fmax(std::fmin(rng.gen_f32(), 0.9999f), 0.0001f);
    }
    std::vector<double> log_probs((size_t)n_cells * k), losses(runs);
    std::vector<uint32_t> iterations(runs);
    uint32_t best = 0;
    CLX(clx_soup_restarts(ctx, k, (uint32_t)runs, centers.data(), log_probs.data(), nullptr, losses.data(), iterations.data(), &best));
    print_results(barcodes, log_probs, n_cells, k, losses[best]);
    clx_destroy(ctx);
    return 0;
}
