// Prometheus HTTP API response -> columnar samples.  HOST code only.
//
// The reference reads Prometheus through `requests.json()` and picks `['data']['result'][0]['value'][1]`, a STRING,
// one node and one query at a time (core/apps/kubeops_api/prometheus_client.py:91-116, :67-72).  The rollup kernels
// want the opposite layout: every series of a `query_range` matrix (or every sample of an instant vector) as one
// `double values[]` column with `int64 offsets[]`.  This file is the step in between: a structural scan of the
// response body that finds the series and counts their samples (ko_prom_open), and a fill pass that converts the
// sample strings straight into caller buffers -- pinned host memory if the caller wants to hand them to
// ko_series_rollup -- on several threads (ko_prom_fill).  Conversion is std::from_chars (correctly rounded, the
// same value Python's float() gives); "NaN", "+Inf", "-Inf" as Prometheus prints them.
//
// Wire format (Prometheus HTTP API v1):
//   {"status":"success","data":{"resultType":"matrix","result":[{"metric":{...},"values":[[t,"v"],...]},...]}}
//   {"status":"success","data":{"resultType":"vector","result":[{"metric":{...},"value":[t,"v"]},...]}}
#include <charconv>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <string>
#include <thread>
#include <vector>

#include "ko_common.cuh"

struct ko_prom_series {
    int64_t values_begin;   // byte offset of the '[' that opens "values" (or "value")
    int64_t n_samples;
    int64_t metric_begin, metric_end;   // byte range of the "metric" object, [begin, end)
    bool single;            // "value": one pair instead of an array of pairs
};

struct ko_prom_doc {
    const char* p = nullptr;
    int64_t len = 0;
    std::vector<ko_prom_series> series;
    int64_t n_samples = 0;
};

namespace {

struct Scanner {
    const char* p;
    int64_t n, i = 0;
    const char* err = nullptr;

    void fail(const char* m) {
        if (!err) err = m;
    }
    void ws() {
        while (i < n && (p[i] == ' ' || p[i] == '\n' || p[i] == '\t' || p[i] == '\r')) ++i;
    }
    bool eat(char c) {
        ws();
        if (i < n && p[i] == c) {
            ++i;
            return true;
        }
        return false;
    }
    // a JSON string; returns the byte range of its contents (escapes left as they are)
    bool str(int64_t* b, int64_t* e) {
        ws();
        if (i >= n || p[i] != '"') {
            fail("expected a string");
            return false;
        }
        ++i;
        *b = i;
        while (i < n && p[i] != '"') i += (p[i] == '\\') ? 2 : 1;
        if (i >= n) {
            fail("unterminated string");
            return false;
        }
        *e = i;
        ++i;
        return true;
    }
    bool key_is(int64_t b, int64_t e, const char* k) const {
        const int64_t l = (int64_t)strlen(k);
        return e - b == l && memcmp(p + b, k, (size_t)l) == 0;
    }
    // skip any JSON value
    void skip() {
        ws();
        if (i >= n) {
            fail("unexpected end of input");
            return;
        }
        const char c = p[i];
        if (c == '"') {
            int64_t b, e;
            str(&b, &e);
        } else if (c == '{' || c == '[') {
            int depth = 0;
            while (i < n) {
                const char d = p[i];
                if (d == '"') {
                    int64_t b, e;
                    if (!str(&b, &e)) return;
                    continue;
                }
                if (d == '{' || d == '[') ++depth;
                if (d == '}' || d == ']') {
                    --depth;
                    if (depth == 0) {
                        ++i;
                        return;
                    }
                }
                ++i;
            }
            fail("unterminated object or array");
        } else {
            while (i < n && p[i] != ',' && p[i] != '}' && p[i] != ']' && p[i] != ' ' && p[i] != '\n') ++i;
        }
    }
};

}  // namespace

namespace {

// "values": [[t,"v"],[t,"v"],...]  -> number of pairs; leaves the scanner behind the closing ']'.  One tight loop
// over the bytes (this is most of the body): strings are skipped, brackets tracked, a pair is a '[' at depth 1.
int64_t count_pairs(Scanner& s) {
    if (!s.eat('[')) {
        s.fail("\"values\" is not an array");
        return 0;
    }
    const char* p = s.p;
    const int64_t n = s.n;
    int64_t i = s.i, cnt = 0;
    int depth = 1;
    while (i < n) {
        const char c = p[i];
        if (c == '"') {
            ++i;
            while (i < n && p[i] != '"') i += (p[i] == '\\') ? 2 : 1;
            ++i;
            continue;
        }
        if (c == '[') {
            if (depth == 1) ++cnt;
            ++depth;
        } else if (c == ']') {
            if (--depth == 0) {
                s.i = i + 1;
                return cnt;
            }
        } else if (c == '{' || c == '}') {
            break;   // not a list of pairs
        }
        ++i;
    }
    s.i = i;
    s.fail("\"values\" is not a closed array of [timestamp, \"value\"] pairs");
    return cnt;
}

// one element of "result"
void scan_series(Scanner& s, ko_prom_doc* doc) {
    if (!s.eat('{')) {
        s.fail("a result element is not an object");
        return;
    }
    ko_prom_series se;
    se.values_begin = -1;
    se.n_samples = 0;
    se.metric_begin = se.metric_end = 0;
    se.single = false;
    if (!s.eat('}')) {
        for (;;) {
            int64_t kb, ke;
            if (!s.str(&kb, &ke)) return;
            if (!s.eat(':')) {
                s.fail("expected ':'");
                return;
            }
            s.ws();
            if (s.key_is(kb, ke, "metric")) {
                se.metric_begin = s.i;
                s.skip();
                se.metric_end = s.i;
            } else if (s.key_is(kb, ke, "values")) {
                se.values_begin = s.i;
                se.n_samples = count_pairs(s);
            } else if (s.key_is(kb, ke, "value")) {
                se.values_begin = s.i;
                se.single = true;
                se.n_samples = 1;
                s.skip();
            } else {
                s.skip();   // e.g. native "histograms": not samples of a float series
            }
            if (s.err) return;
            if (s.eat(',')) continue;
            if (s.eat('}')) break;
            s.fail("expected ',' or '}' in a result element");
            return;
        }
    }
    if (se.values_begin < 0) se.values_begin = 0;   // a series without float samples: zero samples
    doc->n_samples += se.n_samples;
    doc->series.push_back(se);
}

void scan_data(Scanner& s, ko_prom_doc* doc, bool* typed_ok) {
    if (!s.eat('{')) {
        s.fail("\"data\" is not an object");
        return;
    }
    if (s.eat('}')) return;
    for (;;) {
        int64_t kb, ke;
        if (!s.str(&kb, &ke)) return;
        if (!s.eat(':')) {
            s.fail("expected ':'");
            return;
        }
        if (s.key_is(kb, ke, "resultType")) {
            int64_t vb, ve;
            if (!s.str(&vb, &ve)) return;
            *typed_ok = s.key_is(vb, ve, "matrix") || s.key_is(vb, ve, "vector");
            if (!*typed_ok) {
                s.fail("resultType is neither \"matrix\" nor \"vector\"");
                return;
            }
        } else if (s.key_is(kb, ke, "result")) {
            if (!s.eat('[')) {
                s.fail("\"result\" is not an array (scalar / string results carry no series)");
                return;
            }
            if (!s.eat(']')) {
                for (;;) {
                    scan_series(s, doc);
                    if (s.err) return;
                    if (s.eat(',')) continue;
                    if (s.eat(']')) break;
                    s.fail("expected ',' or ']' in \"result\"");
                    return;
                }
            }
        } else {
            s.skip();
        }
        if (s.err) return;
        if (s.eat(',')) continue;
        if (s.eat('}')) return;
        s.fail("expected ',' or '}' in \"data\"");
        return;
    }
}

// Prometheus prints sample values with Go's strconv.FormatFloat(v, 'f', -1, 64): digits, or NaN / +Inf / -Inf.
bool parse_sample_value(const char* b, const char* e, double* out) {
    const int64_t l = e - b;
    if (l == 3 && memcmp(b, "NaN", 3) == 0) {
        *out = std::numeric_limits<double>::quiet_NaN();
        return true;
    }
    if ((l == 4 && memcmp(b, "+Inf", 4) == 0) || (l == 3 && memcmp(b, "Inf", 3) == 0)) {
        *out = std::numeric_limits<double>::infinity();
        return true;
    }
    if (l == 4 && memcmp(b, "-Inf", 4) == 0) {
        *out = -std::numeric_limits<double>::infinity();
        return true;
    }
    if (l > 0 && *b == '+') ++b;   // from_chars takes no leading '+'
    const auto r = std::from_chars(b, e, *out);
    if (r.ec == std::errc::result_out_of_range) {   // as float(): overflow -> inf, underflow -> 0
        const bool neg = (*b == '-');
        bool tiny = false;
        for (const char* q = b; q < e; ++q)
            if ((*q == 'e' || *q == 'E') && q + 1 < e && q[1] == '-') tiny = true;
        *out = tiny ? (neg ? -0.0 : 0.0)
                    : (neg ? -std::numeric_limits<double>::infinity() : std::numeric_limits<double>::infinity());
        return r.ptr == e;
    }
    return r.ec == std::errc() && r.ptr == e;
}

// fill the samples of one series; returns false on a malformed pair
bool fill_series(const ko_prom_doc* doc, const ko_prom_series& se, double* values, double* stamps) {
    Scanner s{doc->p, doc->len};
    s.i = se.values_begin;
    if (se.n_samples == 0) return true;
    if (!se.single && !s.eat('[')) return false;
    for (int64_t k = 0; k < se.n_samples; ++k) {
        if (!s.eat('[')) return false;
        s.ws();
        const int64_t tb = s.i;
        while (s.i < s.n && s.p[s.i] != ',' && s.p[s.i] != ' ') ++s.i;
        if (stamps) {
            double t = 0.0;
            const auto r = std::from_chars(s.p + tb, s.p + s.i, t);
            if (r.ec != std::errc() || r.ptr != s.p + s.i) return false;
            stamps[k] = t;
        }
        if (!s.eat(',')) return false;
        int64_t vb, ve;
        if (!s.str(&vb, &ve)) return false;
        if (!parse_sample_value(s.p + vb, s.p + ve, &values[k])) return false;
        if (!s.eat(']')) return false;
        if (k + 1 < se.n_samples && !s.eat(',')) return false;
    }
    return true;
}

}  // namespace

extern "C" int ko_prom_open(const char* json, int64_t len, ko_prom_doc** out) {
    KO_REQUIRE(json && len >= 0 && out, "ko_prom_open: NULL argument");
    ko_prom_doc* doc = new ko_prom_doc();
    doc->p = json;
    doc->len = len;
    Scanner s{json, len};
    bool typed = false, have_status = false, success = false;
    std::string api_error;
    if (!s.eat('{')) s.fail("the response is not a JSON object");
    if (!s.err && !s.eat('}')) {
        for (;;) {
            int64_t kb, ke;
            if (!s.str(&kb, &ke)) break;
            if (!s.eat(':')) {
                s.fail("expected ':'");
                break;
            }
            if (s.key_is(kb, ke, "status")) {
                int64_t vb, ve;
                if (!s.str(&vb, &ve)) break;
                have_status = true;
                success = s.key_is(vb, ve, "success");
            } else if (s.key_is(kb, ke, "error")) {
                int64_t vb, ve;
                if (!s.str(&vb, &ve)) break;
                api_error.assign(json + vb, (size_t)(ve - vb));
            } else if (s.key_is(kb, ke, "data")) {
                scan_data(s, doc, &typed);
            } else {
                s.skip();
            }
            if (s.err) break;
            if (s.eat(',')) continue;
            if (s.eat('}')) break;
            s.fail("expected ',' or '}'");
            break;
        }
    }
    if (s.err) {
        ko_set_error("ko_prom_open: malformed response near byte %lld: %s", (long long)s.i, s.err);
        delete doc;
        return KO_EINVAL;
    }
    // the reference's guard: `if json['status'] == 'success'` (prometheus_client.py:94)
    if (!have_status || !success) {
        ko_set_error("ko_prom_open: status is not \"success\"%s%s", api_error.empty() ? "" : ": ", api_error.c_str());
        delete doc;
        return KO_ESTATE;
    }
    *out = doc;
    return KO_OK;
}

extern "C" int ko_prom_dims(const ko_prom_doc* doc, int64_t* n_series, int64_t* n_samples) {
    KO_REQUIRE(doc && n_series && n_samples, "ko_prom_dims: NULL argument");
    *n_series = (int64_t)doc->series.size();
    *n_samples = doc->n_samples;
    return KO_OK;
}

extern "C" int ko_prom_fill(const ko_prom_doc* doc, double* values, double* timestamps, int64_t* offsets,
                            int64_t* metric_begin, int64_t* metric_end, int n_threads) {
    KO_REQUIRE(doc && offsets, "ko_prom_fill: NULL argument");
    KO_REQUIRE(values || doc->n_samples == 0, "ko_prom_fill: values is NULL");
    const int64_t n = (int64_t)doc->series.size();
    int64_t acc = 0;
    for (int64_t i = 0; i < n; ++i) {
        offsets[i] = acc;
        acc += doc->series[i].n_samples;
        if (metric_begin) metric_begin[i] = doc->series[i].metric_begin;
        if (metric_end) metric_end[i] = doc->series[i].metric_end;
    }
    offsets[n] = acc;
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 64) n_threads = 64;
    if ((int64_t)n_threads > n) n_threads = (int)(n > 0 ? n : 1);
    // contiguous ranges of series with about the same number of samples each
    std::vector<int64_t> cut(n_threads + 1, n);
    cut[0] = 0;
    for (int t = 1, i = 0; t < n_threads; ++t) {
        const int64_t want = acc * t / n_threads;
        while (i < n && offsets[i] < want) ++i;
        cut[t] = i;
    }
    std::vector<int> bad(n_threads, 0);
    auto work = [&](int t) {
        for (int64_t i = cut[t]; i < cut[t + 1]; ++i) {
            const ko_prom_series& se = doc->series[i];
            if (!fill_series(doc, se, values + offsets[i], timestamps ? timestamps + offsets[i] : nullptr)) {
                bad[t] = 1;
                return;
            }
        }
    };
    if (n_threads == 1) {
        work(0);
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < n_threads; ++t) th.emplace_back(work, t);
        for (auto& x : th) x.join();
    }
    for (int t = 0; t < n_threads; ++t)
        if (bad[t]) {
            ko_set_error("ko_prom_fill: a sample is not a [number, \"number\"] pair");
            return KO_EINVAL;
        }
    return KO_OK;
}

extern "C" void ko_prom_close(ko_prom_doc* doc) { delete doc; }
