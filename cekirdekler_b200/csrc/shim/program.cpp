// program.cpp — kernel-source front-end of libKutuphaneCL.so: source rewrite + NVRTC → sm_100a cubin.
//
// Replaces the clBuildProgram step behind createProgram (reference: ClProgram.cs:64-67 called from
// Worker.cs:261-272).  The reference passes the user's kernel string unmodified to the OpenCL device
// compiler; here the string is (1) rewritten so that every `__kernel void f(args)` / `kernel void f(args)`
// (the spellings ClNumberCruncher's regex recognises, ClNumberCruncher.cs:219-228) becomes
// `extern "C" __global__ void f(const unsigned int __ck_goff, args)`, (2) prefixed with ck_prelude.cuh and
// (3) compiled by NVRTC straight to a cubin for the device's architecture (sm_100a on B200).
// Compiled images are cached per (source, arch) so that the N workers of one cruncher compile once.
#include "handles.h"

#include <nvrtc.h>

#include <cctype>
#include <cstring>
#include <map>
#include <sstream>

namespace ck {

static const char* kPrelude =
#include "ck_prelude_embed.inc"
    ;

static bool ident_char(char c) { return std::isalnum((unsigned char)c) || c == '_'; }

// If src[i..] reads `( <vector type> ) (` returns the length of that prefix (through the second parenthesis), else 0.
static size_t vector_literal_prefix(const std::string& src, size_t i) {
    static const char* bases[] = {"float", "double", "int", "uint", "char", "uchar", "short", "ushort"};
    const size_t n = src.size();
    size_t p = i + 1;
    while (p < n && std::isspace((unsigned char)src[p])) p++;
    size_t b = p;
    while (p < n && ident_char(src[p])) p++;
    if (p - b < 4 || p >= n) return 0;
    const char width = src[p - 1];
    if (width != '2' && width != '3' && width != '4') return 0;
    bool known = false;
    for (const char* base : bases) known = known || src.compare(b, p - 1 - b, base) == 0;
    if (!known) return 0;
    while (p < n && std::isspace((unsigned char)src[p])) p++;
    if (p >= n || src[p] != ')') return 0;
    p++;
    while (p < n && std::isspace((unsigned char)src[p])) p++;
    if (p >= n || src[p] != '(') return 0;
    return p + 1 - i;
}

// Finds `kernel void name(` / `__kernel void name(` occurrences outside comments and rewrites them.
std::string rewrite_kernels(const std::string& src, std::vector<std::pair<std::string, int>>* kernels_out) {
    std::string out;
    out.reserve(src.size() + 256);
    const size_t n = src.size();
    size_t i = 0;
    while (i < n) {
        // copy comments and string literals verbatim
        if (src[i] == '/' && i + 1 < n && src[i + 1] == '/') {
            size_t e = src.find('\n', i);
            if (e == std::string::npos) e = n;
            out.append(src, i, e - i);
            i = e;
            continue;
        }
        if (src[i] == '/' && i + 1 < n && src[i + 1] == '*') {
            size_t e = src.find("*/", i + 2);
            e = (e == std::string::npos) ? n : e + 2;
            out.append(src, i, e - i);
            i = e;
            continue;
        }
        if (src[i] == '"') {
            size_t e = i + 1;
            while (e < n && src[e] != '"') e += (src[e] == '\\') ? 2 : 1;
            e = (e < n) ? e + 1 : n;
            out.append(src, i, e - i);
            i = e;
            continue;
        }
        // OpenCL C vector literal `(float4)(a, b, c, d)` → `make_float4(a, b, c, d)`; the prelude adds the broadcast and
        // mixed scalar/vector overloads OpenCL allows.  `(float4)x` (no second parenthesis) is left alone.
        if (src[i] == '(') {
            size_t len = vector_literal_prefix(src, i);
            if (len) {
                size_t b = i + 1;
                while (std::isspace((unsigned char)src[b])) b++;
                size_t e = b;
                while (ident_char(src[e])) e++;
                out += "make_";
                out.append(src, b, e - b);
                out.push_back('(');
                i += len;
                continue;
            }
        }
        // OpenCL C component selections CUDA's vector structs lack: `.s0`..`.s3` become `.x`..`.w`; `.xy .zw .lo .hi .xyz` wrap
        // the postfix expression they follow into ck_xy(...) etc., which the prelude defines for the float/int/uint vectors
        // (readable and assignable) and, as a fall-back, for any struct that has a member of that name.
        bool after_operand = src[i] == '.' && i > 0 && i + 2 < n && (ident_char(src[i - 1]) || src[i - 1] == ')' || src[i - 1] == ']');
        if (after_operand && ident_char(src[i - 1])) {   // `1.5f`, `2.` are numbers, not operands
            size_t k = i;
            while (k > 0 && ident_char(src[k - 1])) k--;
            if (std::isdigit((unsigned char)src[k])) after_operand = false;
        }
        if (after_operand) {
            size_t e = i + 1;
            while (e < n && ident_char(src[e])) e++;
            const std::string word = src.substr(i + 1, e - i - 1);
            if (word.size() == 2 && word[0] == 's' && word[1] >= '0' && word[1] <= '3') {
                out.push_back('.');
                out.push_back("xyzw"[word[1] - '0']);
                i = e;
                continue;
            }
            if (word == "xy" || word == "zw" || word == "lo" || word == "hi" || word == "xyz") {
                // walk back over the postfix expression already copied to `out`: names, a.b, a->b, f(...), a[...]
                size_t j = out.size();
                for (;;) {
                    if (j > 0 && (out[j - 1] == ')' || out[j - 1] == ']')) {
                        const char close = out[j - 1], open = close == ')' ? '(' : '[';
                        int depth = 0;
                        size_t k = j;
                        while (k > 0) {
                            k--;
                            if (out[k] == close) depth++;
                            else if (out[k] == open && --depth == 0) break;
                        }
                        if (depth != 0) break;
                        j = k;
                        continue;
                    }
                    size_t k = j;
                    while (k > 0 && ident_char(out[k - 1])) k--;
                    if (k == j) break;
                    j = k;
                    if (j > 0 && out[j - 1] == '.') { j--; continue; }
                    if (j > 1 && out[j - 1] == '>' && out[j - 2] == '-') { j -= 2; continue; }
                    break;
                }
                if (j < out.size()) {
                    out.insert(j, "ck_" + word + "(");
                    out.push_back(')');
                    i = e;
                    continue;
                }
            }
        }
        // candidate token start?
        bool at_token_start = (i == 0) || !ident_char(src[i - 1]);
        size_t tok = 0;
        if (at_token_start) {
            if (src.compare(i, 8, "__kernel") == 0 && (i + 8 >= n || !ident_char(src[i + 8]))) tok = 8;
            else if (src.compare(i, 6, "kernel") == 0 && (i + 6 >= n || !ident_char(src[i + 6]))) tok = 6;
        }
        if (!tok) {
            out.push_back(src[i++]);
            continue;
        }
        // kernel \s+ void \s+ name \s* (
        size_t p = i + tok;
        size_t q = p;
        while (q < n && std::isspace((unsigned char)src[q])) q++;
        if (q == p || src.compare(q, 4, "void") != 0 || (q + 4 < n && ident_char(src[q + 4]))) {
            out.append(src, i, tok);
            i += tok;
            continue;
        }
        size_t r = q + 4;
        size_t r0 = r;
        while (r < n && std::isspace((unsigned char)src[r])) r++;
        size_t name_b = r;
        while (r < n && ident_char(src[r])) r++;
        size_t name_e = r;
        while (r < n && std::isspace((unsigned char)src[r])) r++;
        if (r0 == name_b || name_e == name_b || r >= n || src[r] != '(') {
            out.append(src, i, tok);
            i += tok;
            continue;
        }
        // matching ')'
        size_t depth = 0, close = std::string::npos;
        for (size_t k = r; k < n; k++) {
            if (src[k] == '(') depth++;
            else if (src[k] == ')') {
                if (--depth == 0) { close = k; break; }
            }
        }
        if (close == std::string::npos) {
            out.append(src, i, tok);
            i += tok;
            continue;
        }
        std::string name = src.substr(name_b, name_e - name_b);
        std::string params = src.substr(r + 1, close - r - 1);
        // count parameters (commas at depth 0); "(void)" and "()" mean none
        std::string trimmed;
        for (char c : params) if (!std::isspace((unsigned char)c)) trimmed.push_back(c);
        int count = 0;
        if (!trimmed.empty() && trimmed != "void") {
            count = 1;
            int d = 0;
            for (char c : params) {
                if (c == '(' || c == '[' || c == '<') d++;
                else if (c == ')' || c == ']' || c == '>') d--;
                else if (c == ',' && d == 0) count++;
            }
        }
        out += "extern \"C\" __global__ void ";
        {
            // `// ck-launch-bounds: <kernel> <max threads> [<min work-groups per SM>]` anywhere in the source: the attribute
            // cannot sit inside `kernel void name(`, which is what the reference's kernel-name regex must keep matching
            const std::string key = "ck-launch-bounds:";
            for (size_t at = src.find(key); at != std::string::npos; at = src.find(key, at + 1)) {
                std::istringstream line(src.substr(at + key.size(), src.find('\n', at) - at - key.size()));
                std::string kname;
                int threads = 0, blocks = 0;
                line >> kname >> threads;
                if (kname != name || threads <= 0) continue;
                line >> blocks;
                out += "__launch_bounds__(" + std::to_string(threads) + (blocks > 0 ? ", " + std::to_string(blocks) : std::string()) + ") ";
                break;
            }
        }
        out += name;
        out += "(const unsigned int __ck_goff";
        if (count > 0) {
            out += ", ";
            out += params;
        }
        out += ")";
        if (kernels_out) kernels_out->emplace_back(name, count);
        i = close + 1;
    }
    return out;
}

// `// ck-nvrtc: <option> <option>` lines let a kernel string choose compiler flags (the reference API has
// no other channel than the source text: F9 in SURVEY.md).
static void collect_options(const std::string& src, std::vector<std::string>* opts) {
    const std::string key = "ck-nvrtc:";
    size_t pos = 0;
    while ((pos = src.find(key, pos)) != std::string::npos) {
        size_t e = src.find('\n', pos);
        if (e == std::string::npos) e = src.size();
        std::istringstream is(src.substr(pos + key.size(), e - pos - key.size()));
        std::string o;
        while (is >> o) opts->push_back(o);
        pos = e;
    }
}

static std::mutex g_cache_mu;
static std::map<std::string, std::shared_ptr<Cubin>> g_cache;

std::shared_ptr<Cubin> compile_source(const std::string& user_src, int cc_major, int cc_minor) {
    char arch[64];
    // family-specific ("a") targets exist from sm_90 on; B200 is 10.0 → sm_100a
    snprintf(arch, sizeof arch, "--gpu-architecture=sm_%d%d%s", cc_major, cc_minor, cc_major >= 9 ? "a" : "");
    std::string key = std::string(arch) + "\n" + user_src;
    {
        std::lock_guard<std::mutex> lk(g_cache_mu);
        auto it = g_cache.find(key);
        if (it != g_cache.end()) return it->second;
    }
    auto cb = std::make_shared<Cubin>();
    // OpenCL 2.0 device-side enqueue (the reference looks for this very token to make a default device queue,
    // ClNumberCruncher.cs:203-205, Worker.cs:224-227) has no translation here: CUDA's counterpart is dynamic parallelism
    // (`child<<<grid, block>>>(...)` inside a kernel), which needs relocatable device code linked against cudadevrt, and the
    // blocks/ndrange_t argument forms do not map onto it textually.  Say so instead of failing on an undefined identifier.
    {
        const size_t at = user_src.find("enqueue_kernel");
        if (at != std::string::npos && (at == 0 || !ident_char(user_src[at - 1])) && user_src.find('(', at) != std::string::npos) {
            cb->err = 2000;
            cb->log = "Kernel source uses enqueue_kernel(...): OpenCL 2.0 device-side enqueue is not supported by libKutuphaneCL.so on CUDA.\n"
                      "Launch the child work from the host as a second kernel name of the same compute() call (\"parent child\"), "
                      "or restructure the parent to loop over the child's range.";
            return cb;
        }
    }
    std::string body = rewrite_kernels(user_src, &cb->kernels);
    std::string full = std::string(kPrelude) + "\n#line 1 \"user_kernels.cu\"\n" + body;

    std::vector<std::string> opts = {arch, "-lineinfo", "--std=c++17", "-default-device"};
    collect_options(user_src, &opts);
    std::vector<const char*> copts;
    for (auto& o : opts) copts.push_back(o.c_str());

    nvrtcProgram prog = nullptr;
    nvrtcResult r = nvrtcCreateProgram(&prog, full.c_str(), "ck_program.cu", 0, nullptr, nullptr);
    if (r != NVRTC_SUCCESS) {
        cb->err = 1000 + (int)r;
        cb->log = std::string("nvrtcCreateProgram: ") + nvrtcGetErrorString(r);
        return cb;
    }
    r = nvrtcCompileProgram(prog, (int)copts.size(), copts.data());
    size_t logsz = 0;
    nvrtcGetProgramLogSize(prog, &logsz);
    if (logsz > 1) {
        cb->log.resize(logsz);
        nvrtcGetProgramLog(prog, &cb->log[0]);
        while (!cb->log.empty() && cb->log.back() == '\0') cb->log.pop_back();
    }
    if (r != NVRTC_SUCCESS) {
        cb->err = 1000 + (int)r;
        cb->log = std::string("NVRTC build failed (") + nvrtcGetErrorString(r) + ") for " + arch + "\n" + cb->log;
        nvrtcDestroyProgram(&prog);
        return cb;  // failures are not cached: the caller may fix the source
    }
    size_t sz = 0;
    r = nvrtcGetCUBINSize(prog, &sz);
    if (r == NVRTC_SUCCESS && sz > 0) {
        cb->image.resize(sz);
        r = nvrtcGetCUBIN(prog, cb->image.data());
    }
    if (r != NVRTC_SUCCESS || sz == 0) {
        cb->err = 1000 + (int)r;
        cb->log += "\nnvrtcGetCUBIN failed";
    }
    nvrtcDestroyProgram(&prog);
    if (cb->err == 0) {
        std::lock_guard<std::mutex> lk(g_cache_mu);
        g_cache[key] = cb;
    }
    return cb;
}

}  // namespace ck

// Test hooks (exported, no GPU needed): lets the CPU test-suite check the source rewrite and that NVRTC can
// build the shipped kernels for sm_100a without a device present.
extern "C" {
// returns a malloc'd, NUL-terminated rewritten source; caller frees with free()
char* ckRewriteKernelSource(const char* src) {
    std::string out = ck::rewrite_kernels(src ? src : "", nullptr);
    char* p = (char*)malloc(out.size() + 1);
    memcpy(p, out.c_str(), out.size() + 1);
    return p;
}
// compiles `src` for sm_<major><minor>; returns 0 on success; *log_out (malloc'd) holds the build log
int ckCompileCheck(const char* src, int cc_major, int cc_minor, char** log_out, long long* cubin_bytes) {
    auto cb = ck::compile_source(src ? src : "", cc_major, cc_minor);
    if (log_out) {
        *log_out = (char*)malloc(cb->log.size() + 1);
        memcpy(*log_out, cb->log.c_str(), cb->log.size() + 1);
    }
    if (cubin_bytes) *cubin_bytes = (long long)cb->image.size();
    return cb->err;
}
// copies the cached cubin image for inspection with cuobjdump; returns bytes written (0 if absent/too small)
long long ckCompileGetCubin(const char* src, int cc_major, int cc_minor, void* dst, long long cap) {
    auto cb = ck::compile_source(src ? src : "", cc_major, cc_minor);
    if (cb->err || (long long)cb->image.size() > cap) return 0;
    memcpy(dst, cb->image.data(), cb->image.size());
    return (long long)cb->image.size();
}
}
