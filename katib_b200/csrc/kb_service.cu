// kb_service.cu -- host-side mirror of Katib's goptuna suggestion service (plain C++; no device code in this file).
//
// The reference service is Go (pkg/suggestion/v1beta1/goptuna); Go is not available in this image, so the service
// logic that sits above the C ABI is written here in C++, function for function, and speaks the same wire format:
// requests and replies cross kb_service_* as serialized protobuf (pkg/apis/manager/v1beta1/api.proto), which is what
// a 30-line Go main or any gRPC front end hands over unchanged.
//
//   SuggestionService.GetSuggestions              service.go:54-93      -> Service::GetSuggestions
//   initStudyAndSearchSpaceAtFirstRun             service.go:153-171    -> Service::InitAtFirstRun
//   syncTrials                                    service.go:96-151     -> Service::SyncTrials
//   ValidateAlgorithmSettings                     service.go:173-211    -> Service::Validate
//   toGoptunaSampler                              converter.go:38-108   -> ParseAlgorithm
//   toGoptunaSearchSpace                          converter.go:110-184  -> ParseSearchSpace
//   toGoptunaState / getFinalMetric / toGoptunaTrials / toGoptunaParams   converter.go:186-343 -> ConvertTrials
//   sampleNextParam                               sample.go:28-95       -> Service::SampleNext
//   findGoptunaTrialIDByParam                     sample.go:97-127      -> Service::FindTrialByParams
// gRPC codes used by the reference on this path: OK = 0, InvalidArgument = 3, Internal = 13.
#include <algorithm>
#include <charconv>
#include <cmath>
#include <cstdlib>
#include <map>
#include <set>
#include <thread>
#include <unordered_set>

#include "kb_internal.cuh"

enum { GRPC_OK = 0, GRPC_INVALID_ARGUMENT = 3, GRPC_INTERNAL = 13 };
#define KB_SERVICE_SHARD_MIN 65536  // = KB_SUGGEST_SHARD_MIN of kb_tpe_suggest: from here on n_ei_candidates is sharded

// ---------------------------------------------------------------------------------------------------------------
// minimal protobuf wire reader / writer (varint, 64-bit, length-delimited, 32-bit)
// ---------------------------------------------------------------------------------------------------------------
struct PbReader {
    const uint8_t *p, *end;
    bool ok = true;
    PbReader(const void *d, size_t n) : p((const uint8_t *)d), end((const uint8_t *)d + n) {}
    bool done() const { return p >= end || !ok; }
    uint64_t varint() {
        uint64_t v = 0;
        for (int shift = 0; shift < 64 && p < end; shift += 7) {
            uint8_t b = *p++;
            v |= (uint64_t)(b & 0x7F) << shift;
            if (!(b & 0x80)) return v;
        }
        ok = false;
        return 0;
    }
    // returns field number, sets wire type; for length-delimited sets sub-span
    int next(int &wt, PbReader *sub, uint64_t *val) {
        uint64_t key = varint();
        if (!ok) return 0;
        wt = (int)(key & 7);
        const int field = (int)(key >> 3);
        switch (wt) {
            case 0: { uint64_t v = varint(); if (val) *val = v; break; }
            case 1: { if (end - p < 8) { ok = false; return 0; } uint64_t v; memcpy(&v, p, 8); p += 8; if (val) *val = v; break; }
            case 2: {
                uint64_t n = varint();
                if (!ok || (uint64_t)(end - p) < n) { ok = false; return 0; }
                if (sub) *sub = PbReader(p, (size_t)n);
                p += n;
                break;
            }
            case 5: { if (end - p < 4) { ok = false; return 0; } uint32_t v; memcpy(&v, p, 4); p += 4; if (val) *val = v; break; }
            default: ok = false; return 0;
        }
        return field;
    }
    std::string str() const { return std::string((const char *)p, (size_t)(end - p)); }
};
struct PbWriter {
    std::string out;
    void varint(uint64_t v) { while (v >= 0x80) { out.push_back((char)(v | 0x80)); v >>= 7; } out.push_back((char)v); }
    void bytes(int field, const std::string &s) { varint(((uint64_t)field << 3) | 2); varint(s.size()); out += s; }
};

// ---------------------------------------------------------------------------------------------------------------
// messages (only the fields this path reads)
// ---------------------------------------------------------------------------------------------------------------
struct KV { std::string name, value; };
struct ParameterSpecMsg { std::string name; int type = 0; std::string max, min, step; std::vector<std::string> list; };
struct ExperimentMsg {
    std::string name, algorithm_name, objective_metric_name;
    std::vector<KV> settings;
    int objective_type = 0;
    std::vector<ParameterSpecMsg> params;
    bool has_spec = false;
};
struct TrialMsg {
    std::string name, start_time, completion_time;
    uint64_t digest = 0;  // of the serialized message (see digest_bytes)
    int condition = 0;
    std::vector<KV> assignments, metrics;
};
struct GetSuggestionsRequestMsg { ExperimentMsg experiment; std::vector<TrialMsg> trials; int current_request_number = 0, total_request_number = 0; };

static KV parse_kv(PbReader r) {
    KV kv; int wt; PbReader sub(nullptr, 0);
    while (!r.done()) { int f = r.next(wt, &sub, nullptr); if (f == 1 && wt == 2) kv.name = sub.str(); else if (f == 2 && wt == 2) kv.value = sub.str(); }
    return kv;
}
static ParameterSpecMsg parse_param(PbReader r) {
    ParameterSpecMsg ps; int wt; PbReader sub(nullptr, 0); uint64_t v;
    while (!r.done()) {
        int f = r.next(wt, &sub, &v);
        if (f == 1 && wt == 2) ps.name = sub.str();
        else if (f == 2 && wt == 0) ps.type = (int)v;
        else if (f == 3 && wt == 2) {
            PbReader fs = sub; PbReader s2(nullptr, 0); int w2;
            while (!fs.done()) {
                int g = fs.next(w2, &s2, nullptr);
                if (w2 != 2) continue;
                if (g == 1) ps.max = s2.str(); else if (g == 2) ps.min = s2.str(); else if (g == 3) ps.list.push_back(s2.str()); else if (g == 4) ps.step = s2.str();
            }
        }
    }
    return ps;
}
static ExperimentMsg parse_experiment(PbReader r) {
    ExperimentMsg e; int wt; PbReader sub(nullptr, 0); uint64_t v;
    while (!r.done()) {
        int f = r.next(wt, &sub, &v);
        if (f == 1 && wt == 2) e.name = sub.str();
        else if (f == 2 && wt == 2) {  // ExperimentSpec
            e.has_spec = true;
            PbReader spec = sub; PbReader s2(nullptr, 0); int w2; uint64_t v2;
            while (!spec.done()) {
                int g = spec.next(w2, &s2, &v2);
                if (g == 1 && w2 == 2) {  // ParameterSpecs
                    PbReader pss = s2; PbReader s3(nullptr, 0); int w3;
                    while (!pss.done()) { int h = pss.next(w3, &s3, nullptr); if (h == 1 && w3 == 2) e.params.push_back(parse_param(s3)); }
                } else if (g == 2 && w2 == 2) {  // ObjectiveSpec
                    PbReader o = s2; PbReader s3(nullptr, 0); int w3; uint64_t v3;
                    while (!o.done()) { int h = o.next(w3, &s3, &v3); if (h == 1 && w3 == 0) e.objective_type = (int)v3; else if (h == 3 && w3 == 2) e.objective_metric_name = s3.str(); }
                } else if (g == 3 && w2 == 2) {  // AlgorithmSpec
                    PbReader a = s2; PbReader s3(nullptr, 0); int w3;
                    while (!a.done()) { int h = a.next(w3, &s3, nullptr); if (h == 1 && w3 == 2) e.algorithm_name = s3.str(); else if (h == 2 && w3 == 2) e.settings.push_back(parse_kv(s3)); }
                }
            }
        }
    }
    return e;
}
static TrialMsg parse_trial(PbReader r) {
    TrialMsg t; int wt; PbReader sub(nullptr, 0);
    while (!r.done()) {
        int f = r.next(wt, &sub, nullptr);
        if (f == 1 && wt == 2) t.name = sub.str();
        else if (f == 2 && wt == 2) {  // TrialSpec
            PbReader spec = sub; PbReader s2(nullptr, 0); int w2;
            while (!spec.done()) {
                int g = spec.next(w2, &s2, nullptr);
                if (g == 3 && w2 == 2) {
                    PbReader pa = s2; PbReader s3(nullptr, 0); int w3;
                    while (!pa.done()) { int h = pa.next(w3, &s3, nullptr); if (h == 1 && w3 == 2) t.assignments.push_back(parse_kv(s3)); }
                }
            }
        } else if (f == 3 && wt == 2) {  // TrialStatus
            PbReader st = sub; PbReader s2(nullptr, 0); int w2; uint64_t v2;
            while (!st.done()) {
                int g = st.next(w2, &s2, &v2);
                if (g == 1 && w2 == 2) t.start_time = s2.str();
                else if (g == 2 && w2 == 2) t.completion_time = s2.str();
                else if (g == 3 && w2 == 0) t.condition = (int)v2;
                else if (g == 4 && w2 == 2) {
                    PbReader ob = s2; PbReader s3(nullptr, 0); int w3;
                    while (!ob.done()) { int h = ob.next(w3, &s3, nullptr); if (h == 1 && w3 == 2) t.metrics.push_back(parse_kv(s3)); }
                }
            }
        }
    }
    return t;
}
// 64-bit digest of a serialized Trial.  The controller resends EVERY trial with every request
// (suggestionclient.go:111-116); once a trial has reached a finished state here it is never updated again
// (service.go:134-137 pushes value and state once), so a message whose bytes were already seen in that state needs
// neither parsing nor conversion: same bytes, same (already validated) result, and nothing left to do with it.
static uint64_t digest_bytes(const uint8_t *p, size_t n) {
    uint64_t h = 0x9E3779B97F4A7C15ull ^ ((uint64_t)n * 0xFF51AFD7ED558CCDull);
    while (n >= 8) {
        uint64_t v; memcpy(&v, p, 8);
        h = (h ^ v) * 0xFF51AFD7ED558CCDull; h ^= h >> 32;
        p += 8; n -= 8;
    }
    uint64_t v = 0; memcpy(&v, p, n);
    h = (h ^ v) * 0xC4CEB9FE1A85EC53ull; h ^= h >> 29;
    h *= 0xFF51AFD7ED558CCDull; h ^= h >> 32;
    return h;
}

static bool parse_request(const void *data, size_t n, GetSuggestionsRequestMsg &req,
                          const std::unordered_set<uint64_t> *settled = nullptr, long long *skipped = nullptr) {
    PbReader r(data, n); int wt; PbReader sub(nullptr, 0); uint64_t v;
    while (!r.done()) {
        int f = r.next(wt, &sub, &v);
        if (f == 1 && wt == 2) req.experiment = parse_experiment(sub);
        else if (f == 2 && wt == 2) {
            const uint64_t dg = digest_bytes(sub.p, (size_t)(sub.end - sub.p));
            if (settled && settled->count(dg)) { if (skipped) ++*skipped; continue; }
            req.trials.push_back(parse_trial(sub));
            req.trials.back().digest = dg;
        }
        else if (f == 4 && wt == 0) req.current_request_number = (int)(int32_t)v;
        else if (f == 5 && wt == 0) req.total_request_number = (int)(int32_t)v;
    }
    return r.ok;
}

// ---------------------------------------------------------------------------------------------------------------
// Go strconv look-alikes
// ---------------------------------------------------------------------------------------------------------------
static bool go_atoi(const std::string &s, long long &out, std::string &err) {
    const char *b = s.c_str(); char *e = nullptr;
    errno = 0;
    if (s.empty() || isspace((unsigned char)s[0])) { err = "strconv.Atoi: parsing \"" + s + "\": invalid syntax"; return false; }
    long long v = strtoll(b, &e, 10);
    if (e == b || *e != 0 || errno == ERANGE) { err = "strconv.Atoi: parsing \"" + s + "\": invalid syntax"; return false; }
    out = v;
    return true;
}
static bool go_parse_float(const std::string &s, double &out, std::string &err) {
    const char *b = s.c_str(); char *e = nullptr;
    if (s.empty() || isspace((unsigned char)s[0])) { err = "strconv.ParseFloat: parsing \"" + s + "\": invalid syntax"; return false; }
    double v = strtod(b, &e);
    if (e == b || *e != 0) { err = "strconv.ParseFloat: parsing \"" + s + "\": invalid syntax"; return false; }
    out = v;
    return true;
}
// strconv.FormatFloat(p, 'f', -1, 64): shortest digits that round-trip, fixed notation
static std::string go_format_float(double v) {
    char buf[512];
    auto r = std::to_chars(buf, buf + sizeof buf, v, std::chars_format::fixed);
    return std::string(buf, r.ptr);
}
// time.Parse(time.RFC3339Nano, s): YYYY-MM-DDTHH:MM:SS[.frac](Z|+HH:MM|-HH:MM)
static bool rfc3339_ok(const std::string &s) {
    auto dig = [&](size_t i) { return i < s.size() && isdigit((unsigned char)s[i]); };
    if (s.size() < 20) return false;
    for (size_t i : {0u, 1u, 2u, 3u, 5u, 6u, 8u, 9u, 11u, 12u, 14u, 15u, 17u, 18u}) if (!dig(i)) return false;
    if (s[4] != '-' || s[7] != '-' || (s[10] != 'T') || s[13] != ':' || s[16] != ':') return false;
    size_t i = 19;
    if (s[i] == '.') { i++; size_t j = i; while (dig(i)) i++; if (i == j) return false; }
    if (i >= s.size()) return false;
    if (s[i] == 'Z') return i + 1 == s.size();
    if (s[i] != '+' && s[i] != '-') return false;
    return i + 6 == s.size() && dig(i + 1) && dig(i + 2) && s[i + 3] == ':' && dig(i + 4) && dig(i + 5);
}

// ---------------------------------------------------------------------------------------------------------------
// the service
// ---------------------------------------------------------------------------------------------------------------
enum TrialState { ST_RUNNING = 0, ST_COMPLETE = 1, ST_PRUNED = 2, ST_FAIL = 3 };
static bool finished(int st) { return st != ST_RUNNING; }

struct Param {  // one entry of the search space
    std::string name;
    int kind = KB_UNIFORM;
    double low = 0, high = 0, step = 0;
    std::vector<std::string> choices;
};
struct Algo {
    std::string name;
    long long seed = 0; bool has_seed = false;
    double sigma = 0; bool has_sigma = false;
    std::string restart = "none";
    long long n_startup = 10, n_ei = 24;
};
struct InternalTrial {
    int id = 0, state = ST_RUNNING;
    double value = 0;
    std::map<std::string, double> params;  // internal representation by parameter name
    int cma_generation = -1;
    bool cma_told = false;
};
struct ConvertedTrial { std::string name; int state = ST_RUNNING; double value = 0; std::map<std::string, double> params; uint64_t digest = 0; };

struct Status { int code = GRPC_OK; std::string msg; };
static Status fail(int code, const std::string &m) { Status s; s.code = code; s.msg = m; return s; }

struct kb_service {
    std::mutex mu;
    int device = 0;
    kb_ctx *ctx = nullptr;
    // kb_service_create_multi: the other GPUs of this service.  ctx (on devices[0]) is the only context that serves
    // CMA-ES, Sobol, random and small TPE requests; the peers hold replicas of the search space and the TPE history and
    // join in when a TPE request is large enough to shard.  All of them live in this one process, so their winners are
    // merged on the host (kb_tpe_suggest_range): no communicator is needed, and none is made -- NCCL kernels of several
    // GPUs driven from one process can deadlock against a cudaMalloc on another thread, a hazard a merge of D x 24
    // bytes per GPU does not justify.  (One process per GPU -- bench.py, tests/mp_tpe_sharded.py -- uses NCCL.)
    std::vector<int> devices;
    std::vector<kb_ctx *> peers;  // ranks 1 .. n-1
    bool initialised = false;
    Algo algo;
    std::vector<Param> space;
    bool maximize = false;
    std::vector<InternalTrial> trials;
    std::map<std::string, int> mapping;  // Katib trial name -> internal trial id
    // trials not yet claimed by a Katib name, by parameter set, ids ascending.  A trial leaves Running only through
    // SyncTrials, i.e. after it was claimed, so every entry here is Running: the newest match is back().
    std::map<std::map<std::string, double>, std::vector<int>> unclaimed;
    std::vector<int> uploaded;  // trial ids whose rows are on the device as the TPE history, in row order
    long long rows_uploaded = 0, adopted = 0, skipped = 0;
    std::unordered_set<uint64_t> settled;  // digests of Trial messages already synced into a finished state
    bool adopt_unknown = false;  // KB_SVC_ADOPT_UNKNOWN_TRIALS
    bool history_dirty = true;
    uint64_t draw_counter = 0;
    // cmaes
    bool cma_ready = false;
    std::vector<int> cma_dims;  // indices into space of the numeric parameters, sorted by name
    int cma_generation = 0, cma_popsize = 0, cma_popsize0 = 0, cma_opt_generation = 0, cma_funhist_term = 0, cma_restarts = 0;
    int cma_n_large_restarts = 0;
    bool cma_small_regime = true;
    long long cma_small_evals = 0, cma_large_evals = 0;
    double cma_sigma0 = 0;
    std::vector<double> cma_funhist;
    std::string last_message;
    std::string reply;
};

static const char *kFailedStudy = "Failed to create goptuna study and search space: ";

// toGoptunaSampler (converter.go:38-108): settings are parsed, unknown names ignored
static bool ParseAlgorithm(const ExperimentMsg &e, Algo &a, std::string &err) {
    a = Algo();
    a.name = e.algorithm_name;
    for (const KV &s : e.settings) {
        if (a.name == "cmaes") {
            if (s.name == "random_state") { if (!go_atoi(s.value, a.seed, err)) return false; a.has_seed = true; }
            else if (s.name == "sigma") { if (!go_parse_float(s.value, a.sigma, err)) return false; a.has_sigma = true; }
            else if (s.name == "restart_strategy") {
                if (s.value == "ipop" || s.value == "bipop" || s.value == "none") a.restart = s.value;
                else { err = "invalid restart_strategy: '" + s.value + "'"; return false; }
            }
        } else if (a.name == "tpe") {
            if (s.name == "random_state") { if (!go_atoi(s.value, a.seed, err)) return false; a.has_seed = true; }
            else if (s.name == "n_startup_trials") { if (!go_atoi(s.value, a.n_startup, err)) return false; }
            else if (s.name == "n_ei_candidates") { if (!go_atoi(s.value, a.n_ei, err)) return false; }
        } else if (a.name != "sobol") {
            if (s.name == "random_state") { if (!go_atoi(s.value, a.seed, err)) return false; a.has_seed = true; }
        }
    }
    return true;
}

// toGoptunaSearchSpace (converter.go:110-184)
static bool ParseSearchSpace(const ExperimentMsg &e, std::vector<Param> &space, std::string &err) {
    space.clear();
    for (const ParameterSpecMsg &p : e.params) {
        Param q; q.name = p.name;
        if (p.type == 1) {  // DOUBLE
            if (!go_parse_float(p.max, q.high, err) || !go_parse_float(p.min, q.low, err)) return false;
            if (p.step.empty()) q.kind = KB_UNIFORM;
            else { if (!go_parse_float(p.step, q.step, err)) return false; q.kind = KB_DISCRETE; }
        } else if (p.type == 2) {  // INT
            long long hi, lo, st;
            if (!go_atoi(p.max, hi, err) || !go_atoi(p.min, lo, err)) return false;
            q.high = (double)hi; q.low = (double)lo;
            if (p.step.empty()) q.kind = KB_INT;
            else { if (!go_atoi(p.step, st, err)) return false; q.step = (double)st; q.kind = KB_STEP_INT; }
        } else if (p.type == 4 || p.type == 3) {  // CATEGORICAL, and DISCRETE as categorical (converter.go:166-178)
            q.kind = KB_CATEGORICAL; q.choices = p.list;
        } else {
            err = "Unsupported parameter type: " + std::to_string(p.type);
            return false;
        }
        // a Go map keyed by name: a later duplicate replaces the earlier entry
        bool replaced = false;
        for (Param &o : space) if (o.name == q.name) { o = q; replaced = true; }
        if (!replaced) space.push_back(q);
    }
    return true;
}

static Status Validate(const ExperimentMsg *e) {
    if (!e) return fail(GRPC_INVALID_ARGUMENT, "request is empty");
    const std::string &n = e->algorithm_name;
    if (n != "random" && n != "cmaes" && n != "tpe" && n != "sobol") return fail(GRPC_INVALID_ARGUMENT, "unsupported algorithm");
    if (n == "cmaes") {
        int cnt = 0;
        for (const ParameterSpecMsg &p : e->params) if (p.type == 1 || p.type == 2) cnt++;
        if (cnt < 2) return fail(GRPC_INVALID_ARGUMENT, "CMA-ES only supports two or more dimensional continuous search space.");
    }
    std::set<std::string> seen;
    for (const ParameterSpecMsg &p : e->params) {
        if (seen.count(p.name)) return fail(GRPC_INVALID_ARGUMENT, "Detect duplicated parameter name: " + p.name);
        seen.insert(p.name);
    }
    Algo a; std::vector<Param> sp; std::string err;
    if (!ParseAlgorithm(*e, a, err) || !ParseSearchSpace(*e, sp, err)) return fail(GRPC_INTERNAL, kFailedStudy + err);
    return Status();
}

// toGoptunaTrials + toGoptunaParams + toGoptunaState + getFinalMetric (converter.go:186-343)
static bool ConvertTrials(const std::vector<TrialMsg> &in, const std::string &metric, const std::vector<Param> &space,
                          std::vector<ConvertedTrial> &out, std::string &err) {
    out.clear();
    out.reserve(in.size());
    std::map<std::string, size_t> by_name;  // position in `out`
    std::map<std::string, const Param *> param_of;  // first entry of a name wins
    for (const Param &p : space) param_of.emplace(p.name, &p);
    for (const TrialMsg &kt : in) {
        ConvertedTrial t; t.name = kt.name; t.digest = kt.digest;
        if (!kt.start_time.empty() && !rfc3339_ok(kt.start_time)) { err = "parsing time \"" + kt.start_time + "\" as \"2006-01-02T15:04:05.999999999Z07:00\": cannot parse"; return false; }
        if (!kt.completion_time.empty() && !rfc3339_ok(kt.completion_time)) { err = "parsing time \"" + kt.completion_time + "\" as \"2006-01-02T15:04:05.999999999Z07:00\": cannot parse"; return false; }
        switch (kt.condition) {
            case 0: case 1: t.state = ST_RUNNING; break;      // CREATED, RUNNING
            case 2: t.state = ST_COMPLETE; break;             // SUCCEEDED
            case 4: t.state = ST_FAIL; break;                 // FAILED
            case 6: t.state = ST_PRUNED; break;               // EARLYSTOPPED
            default: err = "Unexpected Trial condition: " + std::to_string(kt.condition); return false;
        }
        if (t.state == ST_COMPLETE) {
            bool found = false;
            for (size_t i = kt.metrics.size(); i-- > 0;) {
                if (kt.metrics[i].name != metric) continue;
                if (!go_parse_float(kt.metrics[i].value, t.value, err)) return false;
                found = true;
                break;
            }
            if (!found) { err = "No objective metric in Trial " + kt.name; return false; }
        }
        for (const KV &a : kt.assignments) {
            const auto pit = param_of.find(a.name);
            if (pit == param_of.end()) continue;  // a name outside the search space matches no case of the Go type switch
            const Param *p = pit->second;
            double v = 0;
            if (p->kind == KB_UNIFORM || p->kind == KB_DISCRETE) { if (!go_parse_float(a.value, v, err)) return false; }
            else if (p->kind == KB_INT || p->kind == KB_STEP_INT) {
                long long iv;
                if (!go_atoi(a.value, iv, err)) { err = "strconv.ParseInt: parsing \"" + a.value + "\": invalid syntax"; return false; }
                v = (double)iv;
            } else {
                int idx = -1;
                for (size_t c = 0; c < p->choices.size(); c++) if (p->choices[c] == a.value) { idx = (int)c; break; }
                if (idx < 0) { err = "Invalid categorical value: -1"; return false; }
                v = (double)idx;
            }
            t.params[a.name] = v;
        }
        // a Go map keyed by trial name: a later duplicate replaces the earlier entry
        auto seen = by_name.find(t.name);
        if (seen != by_name.end()) out[seen->second] = t;
        else { by_name.emplace(t.name, out.size()); out.push_back(t); }
    }
    return true;
}

// findGoptunaTrialIDByParam (sample.go:97-127): newest unmapped Running trial with identical parameters
// The reference scans a deep copy of every trial per look-up (O(N) each, O(N^2) for a cold sync); the index keyed by
// the parameter set answers the same question -- same winner -- in O(log N).
static int FindTrialByParams(kb_service *s, const ConvertedTrial &kt) {
    auto it = s->unclaimed.find(kt.params);
    if (it == s->unclaimed.end() || it->second.empty()) return -1;
    const int id = it->second.back();
    it->second.pop_back();
    if (it->second.empty()) s->unclaimed.erase(it);
    return id;
}

static Status SyncTrials(kb_service *s, const std::vector<ConvertedTrial> &kts) {
    for (const ConvertedTrial &kt : kts) {
        int id;
        auto it = s->mapping.find(kt.name);
        if (it == s->mapping.end()) {
            id = FindTrialByParams(s, kt);
            if (id < 0) {
                if (!s->adopt_unknown) return fail(GRPC_INTERNAL, "Same parameter is not found for Trial: " + kt.name);
                // superset behaviour (off by default): a trial this process never suggested -- the pod restarted, or
                // the trials come from a resumed experiment -- joins the study as a Running trial and is then synced
                // like any other; it carries no CMA-ES generation tag, so it never enters a Tell
                InternalTrial t;
                t.id = id = (int)s->trials.size();
                t.params = kt.params;
                s->trials.push_back(t);
                s->adopted++;
            }
            s->mapping[kt.name] = id;
        } else id = it->second;
        InternalTrial &t = s->trials[(size_t)id];
        if (finished(t.state)) { s->settled.insert(kt.digest); continue; }  // finished trials are never updated
        if (kt.state == t.state) continue;
        if (kt.state == ST_COMPLETE) t.value = kt.value;
        t.state = kt.state;
        s->history_dirty = true;
        if (finished(t.state)) s->settled.insert(kt.digest);  // these bytes need no second look (see digest_bytes)
    }
    return Status();
}

static Status engine_error(kb_service *s, const char *what) {
    return fail(GRPC_INTERNAL, std::string(what) + ": " + kb_last_error(s->ctx));
}

// f(rank, ctx) on every context of the service, one host thread per GPU so that the GPUs work at the same time.
// Returns the first non-zero code (and that context, for its error text).
template <class F>
static int on_all_contexts(kb_service *s, F f, kb_ctx **failed = nullptr) {
    const size_t n = 1 + s->peers.size();
    std::vector<int> rc(n, 0);
    std::vector<std::thread> th;
    for (size_t r = 1; r < n; r++) th.emplace_back([&, r] { rc[r] = f((int)r, s->peers[r - 1]); });
    rc[0] = f(0, s->ctx);
    for (auto &t : th) t.join();
    for (size_t r = 0; r < n; r++)
        if (rc[r]) { if (failed) *failed = r == 0 ? s->ctx : s->peers[r - 1]; return rc[r]; }
    return 0;
}

// upload Complete (and value-less Pruned, ranked last) trials as the TPE history
static Status PushHistory(kb_service *s) {
    if (!s->history_dirty) return Status();
    // observations in trial-id order, as study.GetTrials() returns them (the weight ramp and ties depend on the order)
    std::vector<int> ids;
    for (const InternalTrial &t : s->trials) {
        if (t.state != ST_COMPLETE && t.state != ST_PRUNED) continue;
        bool all = true;
        for (const Param &p : s->space) if (!t.params.count(p.name)) all = false;
        if (all) ids.push_back(t.id);
    }
    // Only rows the device does not hold yet cross the boundary -- as long as the rows it holds are still a prefix of
    // the ordered list (trials finishing out of order put a smaller id behind a larger one: then everything is resent).
    size_t keep = 0;
    if (!s->uploaded.empty() && s->uploaded.size() <= ids.size() && std::equal(s->uploaded.begin(), s->uploaded.end(), ids.begin()))
        keep = s->uploaded.size();
    std::vector<double> values, loss;
    for (size_t i = keep; i < ids.size(); i++) {
        const InternalTrial &t = s->trials[(size_t)ids[i]];
        for (const Param &p : s->space) values.push_back(t.params.at(p.name));
        loss.push_back(t.state == ST_COMPLETE ? (s->maximize ? -t.value : t.value) : INFINITY);
    }
    // every GPU of the service builds the same tables from the same rows (no exchange: the build is deterministic)
    const double *vp = values.empty() ? nullptr : values.data(), *lp = loss.empty() ? nullptr : loss.data();
    int rc = 0;
    for (size_t r = 0; r <= s->peers.size() && !rc; r++) {
        kb_ctx *c = r == 0 ? s->ctx : s->peers[r - 1];
        rc = keep > 0 ? kb_tpe_append_history(c, (int64_t)loss.size(), vp, lp) : kb_tpe_set_history(c, (int64_t)loss.size(), vp, lp);
        if (rc) { s->uploaded.clear(); return fail(GRPC_INTERNAL, std::string(keep > 0 ? "kb_tpe_append_history: " : "kb_tpe_set_history: ") + kb_last_error(c)); }
    }
    s->uploaded = ids;
    s->rows_uploaded += (long long)loss.size();
    s->history_dirty = false;
    return Status();
}

static int n_observations(const kb_service *s) {
    int n = 0;
    for (const InternalTrial &t : s->trials) if (t.state == ST_COMPLETE || t.state == ST_PRUNED) n++;
    return n;
}

static uint64_t mix_seed(uint64_t seed, uint64_t counter) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (counter + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// map a unit-cube coordinate to a distribution (sobol): uniform affine; integers / grids by floor; categorical by floor
static double from_unit(const Param &p, double u) {
    switch (p.kind) {
        case KB_UNIFORM: return p.low + u * (p.high - p.low);
        case KB_INT: return std::min(p.high, std::floor(p.low + u * (p.high - p.low + 1.0)));
        case KB_STEP_INT:
        case KB_DISCRETE: { const double nq = std::floor((p.high - p.low) / p.step); return p.low + p.step * std::min(nq, std::floor(u * (nq + 1.0))); }
        default: return std::min((double)p.choices.size() - 1.0, std::floor(u * (double)p.choices.size()));
    }
}

// (re)create the optimiser: cmaes.Sampler's lazy construction, and the restart of the IPOP / BIPOP strategies
static Status CmaInit(kb_service *s, int popsize, bool random_mean) {
    const int n = (int)s->cma_dims.size();
    std::vector<double> mean((size_t)n), lo((size_t)n), hi((size_t)n);
    double sigma0 = INFINITY;
    for (int i = 0; i < n; i++) {
        const Param &p = s->space[(size_t)s->cma_dims[(size_t)i]];
        lo[(size_t)i] = p.low; hi[(size_t)i] = p.high;
        mean[(size_t)i] = p.low + (p.high - p.low) / 2.0;
        if (random_mean) {  // restarts begin from a uniformly drawn point of the box
            const double u = (double)(mix_seed((uint64_t)s->algo.seed ^ 0xC3A5C85C97CB3127ull, s->draw_counter++) >> 11) * (1.0 / 9007199254740992.0);
            mean[(size_t)i] = p.low + u * (p.high - p.low);
        }
        sigma0 = std::min(sigma0, (p.high - p.low) / 6.0);
    }
    if (s->algo.has_sigma) sigma0 = s->algo.sigma;
    if (kb_cmaes_create(s->ctx, n, mean.data(), sigma0, popsize, lo.data(), hi.data())) return engine_error(s, "kb_cmaes_create");
    s->cma_popsize = popsize > 0 ? popsize : 4 + (int)std::floor(3.0 * std::log((double)n));
    s->cma_sigma0 = sigma0;
    s->cma_opt_generation = 0;
    s->cma_funhist_term = 10 + (int)std::ceil(30.0 * n / s->cma_popsize);
    s->cma_funhist.assign((size_t)2 * s->cma_funhist_term, 0.0);
    s->cma_ready = true;
    return Status();
}

// Optimizer.ShouldStop (the `cmaes` library's termination criteria: tolfun over the recent generations, tolx, tolxup,
// no-effect coordinate / axis, condition number)
static bool CmaShouldStop(kb_service *s, Status &st) {
    const int n = (int)s->cma_dims.size();
    std::vector<double> mean((size_t)n), C((size_t)n * n), pc((size_t)n), B((size_t)n * n), D((size_t)n);
    double sigma = 0;
    int gen = 0;
    if (kb_cmaes_get_state(s->ctx, mean.data(), &sigma, C.data(), nullptr, pc.data(), B.data(), D.data(), &gen)) {
        st = engine_error(s, "kb_cmaes_get_state");
        return false;
    }
    const double tolx = 1e-12 * s->cma_sigma0, tolxup = 1e4, tolfun = 1e-12, tolcond = 1e14;
    if (gen > s->cma_funhist_term) {
        const auto mm = std::minmax_element(s->cma_funhist.begin(), s->cma_funhist.end());
        if (*mm.second - *mm.first < tolfun) return true;
    }
    bool all_small = true;
    for (int i = 0; i < n; i++) if (!(sigma * C[(size_t)i * n + i] < tolx && sigma * pc[(size_t)i] < tolx)) all_small = false;
    if (all_small) return true;
    const double dmax = *std::max_element(D.begin(), D.end()), dmin = *std::min_element(D.begin(), D.end());
    if (sigma * dmax > tolxup) return true;
    for (int i = 0; i < n; i++) if (mean[(size_t)i] == mean[(size_t)i] + 0.2 * sigma * std::sqrt(C[(size_t)i * n + i])) return true;
    const int ax = gen % n;
    bool axis_dead = true;
    for (int i = 0; i < n; i++) if (mean[(size_t)i] != mean[(size_t)i] + 0.1 * sigma * D[(size_t)ax] * B[(size_t)i * n + ax]) axis_dead = false;
    if (axis_dead) return true;
    return dmax / dmin > tolcond;
}

static Status CmaStep(kb_service *s, std::vector<double> &x_out) {
    // cmaes.Sampler.SampleRelative: numeric parameters sorted by name; lazily created optimiser; Tell once popsize
    // Complete trials of the current generation exist; restart check; then Ask one solution
    if (!s->cma_ready) {
        s->cma_dims.clear();
        for (size_t i = 0; i < s->space.size(); i++) if (s->space[i].kind != KB_CATEGORICAL) s->cma_dims.push_back((int)i);
        std::sort(s->cma_dims.begin(), s->cma_dims.end(), [&](int a, int b) { return s->space[(size_t)a].name < s->space[(size_t)b].name; });
        Status st = CmaInit(s, 0, false);
        if (st.code) return st;
        s->cma_popsize0 = s->cma_popsize;
        s->cma_generation = 0;
    }
    const int n = (int)s->cma_dims.size();
    std::vector<const InternalTrial *> sol;
    for (const InternalTrial &t : s->trials)
        if (t.state == ST_COMPLETE && t.cma_generation == s->cma_generation && !t.cma_told) sol.push_back(&t);
    if ((int)sol.size() >= s->cma_popsize) {
        std::vector<double> X((size_t)s->cma_popsize * n), f((size_t)s->cma_popsize);
        for (int k = 0; k < s->cma_popsize; k++) {
            for (int i = 0; i < n; i++) X[(size_t)k * n + i] = sol[(size_t)k]->params.at(s->space[(size_t)s->cma_dims[(size_t)i]].name);
            f[(size_t)k] = s->maximize ? -sol[(size_t)k]->value : sol[(size_t)k]->value;
        }
        if (kb_cmaes_tell(s->ctx, X.data(), f.data())) return engine_error(s, "kb_cmaes_tell");
        for (InternalTrial &t : s->trials) if (t.cma_generation == s->cma_generation) t.cma_told = true;
        s->cma_generation += 1;      // the tag written on trials keeps counting across restarts
        s->cma_opt_generation += 1;  // the optimiser's own generation restarts with it
        const auto mm = std::minmax_element(f.begin(), f.end());
        const size_t slot = (size_t)2 * (size_t)(s->cma_opt_generation % s->cma_funhist_term);
        s->cma_funhist[slot] = *mm.first; s->cma_funhist[slot + 1] = *mm.second;
        if (s->algo.restart != "none") {
            Status st;
            const bool stop = CmaShouldStop(s, st);
            if (st.code) return st;
            if (stop) {
                int popsize;
                if (s->algo.restart == "ipop") {
                    popsize = s->cma_popsize * 2;  // cmaes.SamplerOptionIPop(2), converter.go:60
                } else {                          // bipop, cmaes.SamplerOptionBIPop(2), converter.go:62
                    const long long n_eval = (long long)s->cma_popsize * s->cma_opt_generation;
                    (s->cma_small_regime ? s->cma_small_evals : s->cma_large_evals) += n_eval;
                    if (s->cma_small_evals < s->cma_large_evals) {
                        s->cma_small_regime = true;
                        const double mult = std::pow(2.0, (double)s->cma_n_large_restarts);
                        const double u = (double)(mix_seed((uint64_t)s->algo.seed ^ 0x9FB21C651E98DF25ull, s->draw_counter++) >> 11) * (1.0 / 9007199254740992.0);
                        popsize = (int)std::floor(s->cma_popsize0 * std::pow(mult, u * u));
                    } else {
                        s->cma_small_regime = false;
                        s->cma_n_large_restarts += 1;
                        popsize = s->cma_popsize0 * (1 << std::min(s->cma_n_large_restarts, 20));
                    }
                }
                popsize = std::max(2, std::min(popsize, 1 << 20));
                Status st2 = CmaInit(s, popsize, true);
                if (st2.code) return st2;
                s->cma_restarts += 1;
            }
        }
    }
    x_out.assign((size_t)n, 0.0);
    const uint64_t seed = mix_seed((uint64_t)s->algo.seed, s->draw_counter++);
    if (kb_cmaes_ask(s->ctx, 1, nullptr, seed, x_out.data())) return engine_error(s, "kb_cmaes_ask");
    return Status();
}

// sampleNextParam (sample.go:28-95)
static Status SampleNext(kb_service *s, std::vector<KV> &assignments) {
    InternalTrial t;
    t.id = (int)s->trials.size();
    const size_t D = s->space.size();
    std::vector<double> v(D, 0.0);
    const std::string &algo = s->algo.name;
    if (algo == "tpe" && n_observations(s) >= s->algo.n_startup) {
        Status st = PushHistory(s);
        if (st.code) return st;
        const uint64_t seed = mix_seed((uint64_t)s->algo.seed, s->draw_counter++);
        if (!s->peers.empty() && s->algo.n_ei >= KB_SERVICE_SHARD_MIN) {
            // a large candidate set: every GPU draws and scores its index range of the one Philox stream (shard_range of
            // katib_b200/sharding.py); the per-GPU winners are merged here by tpe.Sampler's argmax rule -- best score,
            // lowest candidate index on ties -- which makes the answer that of one GPU scoring the whole set
            const size_t W = 1 + s->peers.size();
            const int64_t n_ei = s->algo.n_ei, base = n_ei / (int64_t)W, extra = n_ei % (int64_t)W;
            std::vector<std::vector<int64_t>> ir(W, std::vector<int64_t>(D, -1));
            std::vector<std::vector<double>> vr(W, std::vector<double>(D, 0.0)), sr(W, std::vector<double>(D, 0.0));
            kb_ctx *bad = nullptr;
            if (on_all_contexts(s, [&](int r, kb_ctx *c) {
                    const int64_t first = r * base + std::min<int64_t>(r, extra), count = base + (r < extra ? 1 : 0);
                    return kb_tpe_suggest_range(c, seed, first, count, ir[(size_t)r].data(), vr[(size_t)r].data(), sr[(size_t)r].data());
                }, &bad))
                return fail(GRPC_INTERNAL, std::string("kb_tpe_suggest: ") + kb_last_error(bad));
            for (size_t d = 0; d < D; d++) {
                int64_t bi = -1; double bs = 0.0;
                for (size_t r = 0; r < W; r++) {
                    if (ir[r][d] < 0) continue;
                    if (bi < 0 || sr[r][d] > bs || (sr[r][d] == bs && ir[r][d] < bi)) { bi = ir[r][d]; bs = sr[r][d]; v[d] = vr[r][d]; }
                }
            }
        } else if (kb_tpe_suggest(s->ctx, s->algo.n_ei, seed, v.data()))
            return engine_error(s, "kb_tpe_suggest");
    } else if (algo == "sobol") {
        std::vector<int> order(D);
        for (size_t i = 0; i < D; i++) order[i] = (int)i;
        std::sort(order.begin(), order.end(), [&](int a, int b) { return s->space[(size_t)a].name < s->space[(size_t)b].name; });
        std::vector<double> u(D);
        const char *skip = getenv("KB_SOBOL_SKIP_FIRST");
        const uint64_t index = s->draw_counter++ + (skip && skip[0] == '1' ? 1 : 0);
        if (kb_sobol_fill(s->ctx, (int)D, index, 1, nullptr, nullptr, u.data())) return engine_error(s, "kb_sobol_fill");
        for (size_t k = 0; k < D; k++) v[(size_t)order[k]] = from_unit(s->space[(size_t)order[k]], u[k]);
    } else {
        // random, TPE's start-up phase, and the independent sampler behind CMA-ES's categorical parameters
        if (kb_random_fill(s->ctx, (uint64_t)s->algo.seed, s->draw_counter++, 1, v.data())) return engine_error(s, "kb_random_fill");
        if (algo == "cmaes") {
            std::vector<double> x;
            Status st = CmaStep(s, x);
            if (st.code) return st;
            for (size_t i = 0; i < s->cma_dims.size(); i++) {
                const Param &p = s->space[(size_t)s->cma_dims[i]];
                double xv = x[i];
                if (p.kind == KB_INT) xv = std::round(xv);
                else if (p.kind == KB_STEP_INT || p.kind == KB_DISCRETE) xv = std::round((xv - p.low) / p.step) * p.step + p.low;
                v[(size_t)s->cma_dims[i]] = std::min(std::max(xv, p.low), p.high);
            }
            t.cma_generation = s->cma_generation;
        }
    }
    assignments.clear();
    for (size_t i = 0; i < D; i++) {
        const Param &p = s->space[i];
        KV a; a.name = p.name;
        if (p.kind == KB_UNIFORM || p.kind == KB_DISCRETE) a.value = go_format_float(v[i]);
        else if (p.kind == KB_INT || p.kind == KB_STEP_INT) a.value = std::to_string((long long)std::llround(v[i]));
        else {
            int c = (int)v[i];
            c = std::max(0, std::min(c, (int)p.choices.size() - 1));
            a.value = p.choices.empty() ? std::string() : p.choices[(size_t)c];
            v[i] = (double)c;
        }
        if (p.kind == KB_INT || p.kind == KB_STEP_INT) v[i] = (double)std::llround(v[i]);
        t.params[p.name] = v[i];
        assignments.push_back(a);
    }
    s->trials.push_back(t);
    s->unclaimed[t.params].push_back(t.id);
    return Status();
}

static Status InitAtFirstRun(kb_service *s, const ExperimentMsg &e) {
    if (s->initialised) return Status();
    std::string err;
    Algo a; std::vector<Param> sp;
    if (!ParseAlgorithm(e, a, err) || !ParseSearchSpace(e, sp, err)) return fail(GRPC_INTERNAL, kFailedStudy + err);
    if (sp.empty()) return fail(GRPC_INTERNAL, std::string(kFailedStudy) + "empty search space");
    if (!s->ctx && kb_create(s->device, &s->ctx)) return fail(GRPC_INTERNAL, std::string(kFailedStudy) + kb_last_error(nullptr));
    if (s->devices.size() > 1 && s->peers.empty()) {
        // one context per further GPU (rank = position in the device list)
        for (size_t r = 1; r < s->devices.size(); r++) {
            kb_ctx *c = nullptr;
            if (kb_create(s->devices[r], &c)) {
                for (kb_ctx *p : s->peers) kb_destroy(p);
                s->peers.clear();
                return fail(GRPC_INTERNAL, std::string(kFailedStudy) + kb_last_error(nullptr));
            }
            s->peers.push_back(c);
        }
    }
    const size_t D = sp.size();
    std::vector<int32_t> kind(D), nch(D);
    std::vector<double> lo(D), hi(D), st(D);
    for (size_t i = 0; i < D; i++) { kind[i] = sp[i].kind; lo[i] = sp[i].low; hi[i] = sp[i].high; st[i] = sp[i].step; nch[i] = (int32_t)sp[i].choices.size(); }
    for (size_t r = 0; r <= s->peers.size(); r++) {
        kb_ctx *c = r == 0 ? s->ctx : s->peers[r - 1];
        if (kb_space_set(c, (int)D, kind.data(), lo.data(), hi.data(), st.data(), nch.data()))
            return fail(GRPC_INTERNAL, std::string(kFailedStudy) + kb_last_error(c));
        kb_tpe_set_option(c, KB_OPT_MERGE_DUPLICATES, 1);  // same mixture, fewer terms (see katib_b200.h)
    }
    s->algo = a; s->space = sp;
    s->maximize = e.objective_type != 1;  // toGoptunaDirection: MINIMIZE -> minimise, anything else -> maximise
    s->initialised = true;
    return Status();
}

static Status GetSuggestions(kb_service *s, const GetSuggestionsRequestMsg &req, std::string &reply) {
    Status st = InitAtFirstRun(s, req.experiment);
    if (st.code) return st;
    std::vector<ConvertedTrial> kts;
    std::string err;
    if (!ConvertTrials(req.trials, req.experiment.objective_metric_name, s->space, kts, err)) return fail(GRPC_INTERNAL, err);
    st = SyncTrials(s, kts);
    if (st.code) return st;
    PbWriter w;
    for (int i = 0; i < req.current_request_number; i++) {
        std::vector<KV> as;
        st = SampleNext(s, as);
        if (st.code) return st;
        PbWriter pa;
        for (const KV &a : as) { PbWriter kv; kv.bytes(1, a.name); kv.bytes(2, a.value); pa.bytes(1, kv.out); }
        w.bytes(1, pa.out);
    }
    reply = w.out;
    return Status();
}

// ---------------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------------
extern "C" int kb_service_create(int device, kb_service **out) {
    if (!out) return KB_ERR_INVALID;
    kb_service *s = new kb_service();
    s->device = device;
    *out = s;
    return KB_OK;
}
extern "C" int kb_service_create_multi(int n_devices, const int *devices, kb_service **out) {
    if (!out) return KB_ERR_INVALID;
    *out = nullptr;
    if (n_devices < 1 || n_devices > 64 || !devices) return kb_fail(nullptr, KB_ERR_INVALID, "kb_service_create_multi: n_devices outside [1, 64] or devices is NULL");
    for (int i = 0; i < n_devices; i++)
        for (int j = 0; j < i; j++)
            if (devices[i] == devices[j]) return kb_fail(nullptr, KB_ERR_INVALID, "kb_service_create_multi: device %d listed twice", devices[i]);
    kb_service *s = new kb_service();
    s->device = devices[0];
    s->devices.assign(devices, devices + n_devices);
    *out = s;
    return KB_OK;
}
extern "C" int kb_service_destroy(kb_service *s) {
    if (!s) return KB_OK;
    for (kb_ctx *p : s->peers) kb_destroy(p);
    if (s->ctx) kb_destroy(s->ctx);
    delete s;
    return KB_OK;
}
extern "C" const char *kb_service_last_message(const kb_service *s) { return s ? s->last_message.c_str() : ""; }

extern "C" int kb_service_get_suggestions(kb_service *s, const void *request, size_t request_len, const void **reply,
                                          size_t *reply_len) {
    if (!s) return GRPC_INTERNAL;
    std::lock_guard<std::mutex> lk(s->mu);
    GetSuggestionsRequestMsg req;
    Status st;
    if (!parse_request(request, request_len, req, &s->settled, &s->skipped)) st = fail(GRPC_INTERNAL, "malformed GetSuggestionsRequest");
    else st = GetSuggestions(s, req, s->reply);
    s->last_message = st.msg;
    if (st.code) s->reply.clear();
    if (reply) *reply = s->reply.data();
    if (reply_len) *reply_len = s->reply.size();
    return st.code;
}

extern "C" int kb_service_validate_algorithm_settings(kb_service *s, const void *request, size_t request_len) {
    if (!s) return GRPC_INTERNAL;
    std::lock_guard<std::mutex> lk(s->mu);
    Status st;
    if (!request) st = Validate(nullptr);
    else {
        PbReader r(request, request_len); int wt; PbReader sub(nullptr, 0);
        ExperimentMsg e;
        while (!r.done()) { int f = r.next(wt, &sub, nullptr); if (f == 1 && wt == 2) e = parse_experiment(sub); }
        st = r.ok ? Validate(&e) : fail(GRPC_INTERNAL, "malformed ValidateAlgorithmSettingsRequest");
    }
    s->last_message = st.msg;
    return st.code;
}

// which: 0 = CMA-ES population size in use, 1 = restarts so far, 2 = internal trials, 3 = generation tag, 4 = Complete observations
extern "C" long long kb_service_stat(kb_service *s, int which) {
    if (!s) return -1;
    std::lock_guard<std::mutex> lk(s->mu);
    switch (which) {
        case 0: return s->cma_popsize;
        case 1: return s->cma_restarts;
        case 2: return (long long)s->trials.size();
        case 3: return s->cma_generation;
        case 4: { long long c = 0; for (const InternalTrial &t : s->trials) c += t.state == ST_COMPLETE; return c; }
        case 5: return s->rows_uploaded;
        case 6: return s->adopted;
        case 7: return s->skipped;
        default: return -1;
    }
}

extern "C" int kb_service_set_option(kb_service *s, int option, int value) {
    if (!s) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(s->mu);
    if (option == KB_SVC_ADOPT_UNKNOWN_TRIALS) { s->adopt_unknown = value != 0; return KB_OK; }
    return KB_ERR_INVALID;
}
