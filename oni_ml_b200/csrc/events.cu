// events.cu -- raw event rows -> (document, word) indices -> scored rows.
//
// Host side of SURVEY.md 8(f) N3: everything flow_post_lda.scala and
// dns_post_lda.scala do around the score loop -- header removal, the field
// count filter, the empirical-CDF cuts, binning, the port -> word rule, the
// sub-domain features, the string lookups into doc_results.csv /
// word_results.csv and the text of the scored rows.  The score loop itself
// (flow_post_lda.scala:232-238, dns_post_lda.scala:315-318), the threshold and
// the ranking run on the GPU through oni_score_flow / oni_score_dns.
//
// Spark is replaced by plain threads over slices of the input; nothing here
// depends on how the slices are cut (the cuts are computed from one sorted copy
// of each column, the lookups are read-only).
//
// Text conventions that live in the JVM and are therefore unpinned by the
// reference: Double.toString is restated as "shortest decimal
// that round-trips" in Java's layout; rows with equal score keep input order.
#include <errno.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <algorithm>
#include <string>
#include <string_view>
#include <thread>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "common.h"

namespace {

using sv = std::string_view;

// ------------------------------------------------------------------ JVM text rules
// String.split(sep) with a one-character separator: trailing empty strings are dropped,
// leading and interior ones kept; "" gives one empty field.
void split_java(sv s, char sep, std::vector<sv> &out)
{
    out.clear();
    size_t a = 0;
    for (;;) {
        const size_t b = s.find(sep, a);
        if (b == sv::npos) { out.push_back(s.substr(a)); break; }
        out.push_back(s.substr(a, b - a));
        a = b + 1;
    }
    while (out.size() > 1 && out.back().empty()) out.pop_back();
    if (out.size() == 1 && out[0].empty() && !s.empty()) out.clear();  // ",,," -> no fields
}

// String.trim: code units <= U+0020 removed from both ends
sv trim_java(sv s)
{
    size_t a = 0, b = s.size();
    while (a < b && (unsigned char)s[a] <= 0x20) a++;
    while (b > a && (unsigned char)s[b - 1] <= 0x20) b--;
    return s.substr(a, b - a);
}

// String.toDouble (Double.parseDouble): surrounding whitespace allowed, the rest must parse
bool parse_double(sv f, double *out)
{
    f = trim_java(f);
    if (f.empty() || f.size() > 63) return false;
    char buf[64];
    memcpy(buf, f.data(), f.size());
    buf[f.size()] = 0;
    char *end = nullptr;
    const double v = strtod(buf, &end);
    if (end == buf) return false;
    if (*end == 'd' || *end == 'D' || *end == 'f' || *end == 'F') end++;  // Java's type suffixes
    if (*end || v != v) return false;  // (a NaN would poison the sort behind the cut vectors)
    *out = v;
    return true;
}

// Double.toString: the shortest digit string that reads back as the same double, written
// d.ddd for 1e-3 <= |v| < 1e7 and d.dddE<n> outside, always with a digit after the point.
int java_double_str(char *out, double v)
{
    if (v != v) return sprintf(out, "NaN");
    if (isinf(v)) return sprintf(out, v < 0 ? "-Infinity" : "Infinity");
    char *p = out;
    if (signbit(v)) { *p++ = '-'; v = -v; }
    if (v == 0.0) { strcpy(p, "0.0"); return (int)(p - out) + 3; }
    if (v < 1e7 && v == (double)(int)v) {  // ports, bins, packet counts: "<integer>.0"
        char tmp[12];
        int n = 0;
        for (int u = (int)v; u > 0; u /= 10) tmp[n++] = (char)('0' + u % 10);
        while (n > 0) *p++ = tmp[--n];
        *p++ = '.'; *p++ = '0'; *p = 0;
        return (int)(p - out);
    }
    char e[40];
    int prec = 0;
    for (prec = 0; prec < 17; prec++) {
        snprintf(e, sizeof e, "%.*e", prec, v);
        if (strtod(e, nullptr) == v) break;
        // Just above a power of two the doubles below lie twice as close as those above, so a
        // decimal of this length that is not the nearest one can still read back as v.
        unsigned long long mant = 0;
        const char *q = e;
        for (; *q && *q != 'e'; q++)
            if (*q >= '0' && *q <= '9') mant = mant * 10 + (unsigned long long)(*q - '0');
        const int ex = atoi(q + 1) - prec;
        char alt[2][40];
        bool ok[2];
        snprintf(alt[0], sizeof alt[0], "%llue%d", mant + 1, ex);
        snprintf(alt[1], sizeof alt[1], "%llue%d", mant - 1, ex);
        for (int a = 0; a < 2; a++) ok[a] = strtod(alt[a], nullptr) == v;
        if (ok[0] || ok[1]) {
            const unsigned long long pick = ok[0] ? mant + 1 : mant - 1;
            char d[24];
            const int nd = snprintf(d, sizeof d, "%llu", pick);
            // back into d.ddde<n> form for the layout code below
            int w = 0;
            e[w++] = d[0];
            e[w++] = '.';
            for (int i = 1; i < nd; i++) e[w++] = d[i];
            snprintf(e + w, sizeof e - w, "e%d", ex + nd - 1);
            break;
        }
    }
    char digits[24];
    int nd = 0;
    const char *q = e;
    for (; *q && *q != 'e'; q++)
        if (*q >= '0' && *q <= '9') digits[nd++] = *q;
    const int e10 = atoi(q + 1);
    while (nd > 1 && digits[nd - 1] == '0') nd--;
    if (v >= 1e-3 && v < 1e7) {
        if (e10 >= 0) {
            for (int i = 0; i <= e10; i++) *p++ = i < nd ? digits[i] : '0';
            *p++ = '.';
            if (nd > e10 + 1) for (int i = e10 + 1; i < nd; i++) *p++ = digits[i];
            else *p++ = '0';
        } else {
            *p++ = '0'; *p++ = '.';
            for (int i = 0; i < -e10 - 1; i++) *p++ = '0';
            for (int i = 0; i < nd; i++) *p++ = digits[i];
        }
    } else {
        *p++ = digits[0]; *p++ = '.';
        if (nd > 1) for (int i = 1; i < nd; i++) *p++ = digits[i];
        else *p++ = '0';
        *p++ = 'E';
        p += sprintf(p, "%d", e10);
    }
    *p = 0;
    return (int)(p - out);
}

// ------------------------------------------------------------------ plumbing
int io_threads()
{
    const char *e = getenv("ONI_IO_THREADS");
    int t = e ? atoi(e) : (int)std::thread::hardware_concurrency();
    return std::max(1, std::min(t, 64));
}

// ONI_LDA_TIMING=1: stage times on stderr
struct Laps {
    bool on = getenv("ONI_LDA_TIMING") != nullptr;
    double t = now();
    static double now()
    {
        struct timespec ts;
        clock_gettime(CLOCK_MONOTONIC, &ts);
        return ts.tv_sec + 1e-9 * ts.tv_nsec;
    }
    void lap(const char *what)
    {
        if (!on) return;
        const double n = now();
        fprintf(stderr, "[oni events] %-28s %8.1f ms\n", what, 1e3 * (n - t));
        t = n;
    }
};

// fn(lo, hi, tid) over [0, n) in contiguous slices; the first slice's message wins on failure
template <class F>
int parallel_rows(size_t n, F fn)
{
    const int T = (int)std::min<size_t>(io_threads(), std::max<size_t>(1, n / 4096));
    std::vector<std::string> err(T);
    std::vector<std::thread> th;
    const size_t per = (n + T - 1) / T;
    for (int t = 0; t < T; t++) {
        const size_t lo = std::min(n, t * per), hi = std::min(n, lo + per);
        th.emplace_back([&, lo, hi, t] { fn(lo, hi, t, err[t]); });
    }
    for (auto &x : th) x.join();
    for (auto &m : err)
        if (!m.empty()) return oni_set_error(ONI_ERR_IO, "%s", m.c_str());
    return ONI_OK;
}

int slurp(const char *path, std::string &buf)
{
    FILE *f = fopen(path, "rb");
    if (!f) return oni_set_error(ONI_ERR_IO, "cannot open %s: %s", path, strerror(errno));
    fseek(f, 0, SEEK_END);
    const long long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    buf.resize((size_t)std::max(0LL, n));
    const size_t got = n > 0 ? fread(&buf[0], 1, (size_t)n, f) : 0;
    fclose(f);
    if ((long long)got != n) return oni_set_error(ONI_ERR_IO, "short read on %s", path);
    return ONI_OK;
}

// sc.textFile records: split at '\n', a '\r' before it dropped, no empty record after the last '\n'
void split_lines(const std::string &buf, std::vector<sv> &lines)
{
    size_t a = 0;
    const size_t n = buf.size();
    while (a < n) {
        const void *nl = memchr(buf.data() + a, '\n', n - a);
        size_t b = nl ? (size_t)((const char *)nl - buf.data()) : n;
        size_t e = b;
        if (e > a && buf[e - 1] == '\r') e--;
        lines.emplace_back(buf.data() + a, e - a);
        a = b + 1;
    }
}

// doc_results.csv / word_results.csv -> key -> row index, values [rows][20]
// (flow_post_lda.scala:105-123): key = split(",")(0), numbers = split(",")(1) without '"',
// split(" "), each toDouble; collectAsMap keeps the last line of a repeated key.
struct Table {
    std::string text;
    std::unordered_map<sv, int32_t> index;
    std::vector<double> values;
    int32_t rows = 0;
};

int load_table(const char *path, Table &t)
{
    ONI_TRY(slurp(path, t.text));
    std::vector<sv> lines, f, nums;
    split_lines(t.text, lines);
    t.index.reserve(lines.size() * 2);
    std::string cleaned;
    for (size_t i = 0; i < lines.size(); i++) {
        split_java(lines[i], ',', f);
        if (f.size() < 2) return oni_set_error(ONI_ERR_IO, "%s line %zu: no ',' separated value list", path, i + 1);
        cleaned.assign(f[1]);
        cleaned.erase(std::remove(cleaned.begin(), cleaned.end(), '"'), cleaned.end());
        split_java(cleaned, ' ', nums);
        if (nums.size() < 20)  // the score loops read elements 0..19
            return oni_set_error(ONI_ERR_IO, "%s line %zu: %zu values, the scorer needs 20", path, i + 1, nums.size());
        auto it = t.index.find(f[0]);
        int32_t row;
        if (it == t.index.end()) {
            row = t.rows++;
            t.index.emplace(f[0], row);
            t.values.resize((size_t)t.rows * 20);
        } else {
            row = it->second;
        }
        for (size_t k = 0; k < nums.size(); k++) {
            double v;
            if (!parse_double(nums[k], &v))
                return oni_set_error(ONI_ERR_IO, "%s line %zu: '%.*s' is not a number", path, i + 1,
                                     (int)nums[k].size(), nums[k].data());
            if (k < 20) t.values[(size_t)row * 20 + k] = v;
        }
    }
    return ONI_OK;
}

int32_t lookup(const Table &t, sv key)
{
    auto it = t.index.find(key);
    return it == t.index.end() ? -1 : it->second;
}

// compute_ecdf + distributed_quantiles (flow_post_lda.scala:67-103): for every quantile q the
// largest sample value whose empirical CDF (count of samples <= value, as a double, divided by
// the sample size) is < q; accumulators start at 0 and only grow (max).
void ecdf_cuts(std::vector<double> &x, const double *quant, int nq, double *cuts)
{
    const int T = io_threads();
    const size_t n = x.size();
    if (T > 1 && n > (1u << 20)) {  // sorted slices, then pairwise merges
        std::vector<size_t> edge(T + 1);
        for (int t = 0; t <= T; t++) edge[t] = n * (size_t)t / T;
        {
            std::vector<std::thread> th;
            for (int t = 0; t < T; t++) th.emplace_back([&, t] { std::sort(x.begin() + edge[t], x.begin() + edge[t + 1]); });
            for (auto &h : th) h.join();
        }
        for (int w = 1; w < T; w *= 2) {
            std::vector<std::thread> th;
            for (int t = 0; t + w < T; t += 2 * w)
                th.emplace_back([&, t, w] {
                    std::inplace_merge(x.begin() + edge[t], x.begin() + edge[t + w], x.begin() + edge[std::min(T, t + 2 * w)]);
                });
            for (auto &h : th) h.join();
        }
    } else {
        std::sort(x.begin(), x.end());
    }
    const double nv = (double)n;
    for (int q = 0; q < nq; q++) {
        cuts[q] = 0.0;
        // e = the largest count with e / n < q  (monotone in e)
        size_t lo = 0, hi = n;  // invariant: lo qualifies (0/n = 0 < q needs q > 0)
        if (!(n > 0 && 0.0 / nv < quant[q])) continue;
        while (lo < hi) {
            const size_t mid = lo + (hi - lo + 1) / 2;
            if ((double)mid / nv < quant[q]) lo = mid; else hi = mid - 1;
        }
        if (lo == 0) continue;
        // the largest distinct value whose run ends at or before position lo
        double v = x[lo - 1];
        if (lo < n && x[lo] == v) {
            const size_t first = std::lower_bound(x.begin(), x.begin() + lo, v) - x.begin();
            if (first == 0) continue;
            v = x[first - 1];
        }
        cuts[q] = std::max(0.0, v);
    }
}

int bin_of(double v, const double *cuts, int n)
{
    int b = 0;
    for (int i = 0; i < n; i++) b += (v > cuts[i]);
    return b;
}

const double QUANT10[10] = {0.0, 0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9};
const double QUINT5[5] = {0, 0.2, 0.4, 0.6, 0.8};

// dns_post_lda.scala:149 -- the script's literal set, "" and "krd" included
const char *const COUNTRY_CODES[] = {
    "ac", "ad", "ae", "af", "ag", "ai", "al", "am", "an", "ao", "aq", "ar", "as", "at", "au", "aw", "ax", "az", "ba",
    "bb", "bd", "be", "bf", "bg", "bh", "bi", "bj", "bm", "bn", "bo", "bq", "br", "bs", "bt", "bv", "bw", "by", "bz",
    "ca", "cc", "cd", "cf", "cg", "ch", "ci", "ck", "cl", "cm", "cn", "co", "cr", "cu", "cv", "cw", "cx", "cy", "cz",
    "de", "dj", "dk", "dm", "do", "dz", "ec", "ee", "eg", "eh", "er", "es", "et", "eu", "fi", "fj", "fk", "fm", "fo",
    "fr", "ga", "gb", "gd", "ge", "gf", "gg", "gh", "gi", "gl", "gm", "gn", "gp", "gq", "gr", "gs", "gt", "gu", "gw",
    "gy", "hk", "hm", "hn", "hr", "ht", "hu", "id", "ie", "il", "im", "in", "io", "iq", "ir", "is", "it", "je", "jm",
    "jo", "jp", "ke", "kg", "kh", "ki", "km", "kn", "kp", "kr", "krd", "kw", "ky", "kz", "la", "lb", "lc", "li", "lk",
    "lr", "ls", "lt", "lu", "lv", "ly", "ma", "mc", "md", "me", "mg", "mh", "mk", "ml", "mm", "mn", "mo", "mp", "mq",
    "mr", "ms", "mt", "mu", "mv", "mw", "mx", "my", "mz", "na", "nc", "ne", "nf", "ng", "ni", "nl", "no", "np", "nr",
    "nu", "nz", "om", "pa", "pe", "pf", "pg", "ph", "pk", "pl", "pm", "pn", "pr", "ps", "pt", "pw", "py", "qa", "re",
    "ro", "rs", "ru", "rw", "sa", "sb", "sc", "sd", "se", "sg", "sh", "si", "sj", "", "sk", "sl", "sm", "sn", "so",
    "sr", "ss", "st", "su", "sv", "sx", "sy", "sz", "tc", "td", "tf", "tg", "th", "tj", "tk", "tl", "tm", "tn", "to",
    "tp", "tr", "tt", "tv", "tw", "tz", "ua", "ug", "uk", "us", "uy", "uz", "va", "vc", "ve", "vg", "vi", "vn", "vu",
    "wf", "ws", "ye", "yt", "za", "zm", "zw"};

}  // namespace

// ------------------------------------------------------------------ the handle
struct oni_events {
    int kind = 0;  // 1 flow, 0 dns
    bool has_tables = false;
    std::string data, top_text;
    Table docs, words;
    int64_t lines_in = 0;
    std::vector<sv> row;  // kept rows (flow: the trimmed line)
    double cuts[5][10] = {};
    int ncuts[5] = {};
    // flow: per-row derived columns
    std::vector<double> num_time, word_port;
    std::vector<uint8_t> bins;  // flow 3 per row (ibyt, ipkt, time); dns 5 per row
    std::vector<uint8_t> flags;  // flow: 1 = pair written "sip dip", 2 = "-1_" on src word, 4 = on dest word
    // dns: per-row derived columns
    std::vector<std::string> domain, subdomain;
    std::vector<int32_t> sub_len, num_parts;
    std::vector<double> entropy;
    std::vector<uint8_t> top;
    std::unordered_set<sv> top_domains;
    std::unordered_set<sv> country;
    // indices handed to the scorer
    std::vector<int32_t> a_doc, b_doc, a_word, b_word;
};

namespace {

// ------------------------------------------------------------------ flow
// the word of one row: word_port _ time_bin _ ibyt_bin _ ipkt_bin, every part a Double's text
// (flow_post_lda.scala:211: the bins went through toDouble at :186-188)
int flow_word(const oni_events &ev, size_t i, char *out)
{
    char *p = out;
    p += java_double_str(p, ev.word_port[i]);
    *p++ = '_';
    p += java_double_str(p, (double)ev.bins[3 * i + 2]);
    *p++ = '_';
    p += java_double_str(p, (double)ev.bins[3 * i + 0]);
    *p++ = '_';
    p += java_double_str(p, (double)ev.bins[3 * i + 1]);
    *p = 0;
    return (int)(p - out);
}

int flow_open(oni_events &ev)
{
    Laps laps;
    std::vector<sv> lines;
    split_lines(ev.data, lines);
    laps.lap("split lines");
    ev.lines_in = (int64_t)lines.size();
    const size_t n = lines.size();
    // removeHeader (:20-24): every line equal to the first one goes
    const sv header = n ? lines[0] : sv();
    std::vector<uint8_t> keep(n, 0);
    std::vector<double> tm(n), ibyt(n), ipkt(n), p10(n), p11(n);
    ONI_TRY(parallel_rows(n, [&](size_t lo, size_t hi, int, std::string &err) {
        std::vector<sv> f;
        for (size_t i = lo; i < hi; i++) {
            if (lines[i] == header) continue;
            split_java(lines[i], ',', f);  // :135 datagood
            if (f.size() != 27) continue;
            const sv t = trim_java(lines[i]);  // :146 _.trim.split(",")
            split_java(t, ',', f);
            if (f.size() != 27) continue;  // trimming emptied the last field: the script's row would be short
            double h, m, s;
            if (!parse_double(f[4], &h) || !parse_double(f[5], &m) || !parse_double(f[6], &s) ||
                !parse_double(f[17], &ibyt[i]) || !parse_double(f[16], &ipkt[i]) || !parse_double(f[10], &p10[i]) ||
                !parse_double(f[11], &p11[i])) {
                if (err.empty()) err = "flow line " + std::to_string(i + 1) + ": a numeric field does not parse";
                return;
            }
            tm[i] = h + m / 60 + s / 3600;  // :141
            lines[i] = t;
            keep[i] = 1;
        }
    }));
    laps.lap("filter + parse");
    size_t m = 0;
    for (size_t i = 0; i < n; i++)
        if (keep[i]) {
            lines[m] = lines[i]; tm[m] = tm[i]; ibyt[m] = ibyt[i]; ipkt[m] = ipkt[i]; p10[m] = p10[i]; p11[m] = p11[i];
            m++;
        }
    lines.resize(m); tm.resize(m); ibyt.resize(m); ipkt.resize(m); p10.resize(m); p11.resize(m);
    ev.row.swap(lines);
    ev.num_time = tm;
    // :148-158 the three cut vectors
    ev.ncuts[0] = 10; ev.ncuts[1] = 10; ev.ncuts[2] = 5;
    {
        std::vector<double> tmp(tm);
        ecdf_cuts(tmp, QUANT10, 10, ev.cuts[0]);
        tmp = ibyt;
        ecdf_cuts(tmp, QUANT10, 10, ev.cuts[1]);
        tmp = ipkt;
        ecdf_cuts(tmp, QUINT5, 5, ev.cuts[2]);
    }
    laps.lap("compact + cuts");
    ev.bins.resize(3 * m); ev.word_port.resize(m); ev.flags.resize(m);
    ev.a_doc.resize(m); ev.b_doc.resize(m); ev.a_word.resize(m); ev.b_word.resize(m);
    const int rc = parallel_rows(m, [&](size_t lo, size_t hi, int, std::string &) {
        std::vector<sv> f;
        char w[160];
        for (size_t i = lo; i < hi; i++) {
            split_java(ev.row[i], ',', f);
            ev.bins[3 * i + 0] = (uint8_t)bin_of(ibyt[i], ev.cuts[1], 10);  // :166-180
            ev.bins[3 * i + 1] = (uint8_t)bin_of(ipkt[i], ev.cuts[2], 5);
            ev.bins[3 * i + 2] = (uint8_t)bin_of(tm[i], ev.cuts[0], 10);
            // adjust_port (:182-218); the script calls column 10 "dport" and column 11 "sport"
            const double dport = p10[i], sport = p11[i];
            const double lo_p = fmin(dport, sport), hi_p = fmax(dport, sport);
            double word_port;
            int p_case;
            if ((dport <= 1024 || sport <= 1024) && (dport > 1024 || sport > 1024) && lo_p != 0) {
                p_case = 2; word_port = lo_p;
            } else if (dport > 1024 && sport > 1024) {
                p_case = 3; word_port = 333333;
            } else if (dport == 0 && sport != 0) {
                p_case = 4; word_port = sport;
            } else if (sport == 0 && dport != 0) {
                p_case = 4; word_port = dport;
            } else {
                p_case = 1;
                word_port = (lo_p == 0) ? hi_p : 111111;
            }
            uint8_t fl = 0;
            if (f[8] < f[9]) fl |= 1;  // `sip != 0` compares a String with an Int: always true
            if (p_case == 2 && dport < sport) fl |= 4;
            else if (p_case == 2 && sport < dport) fl |= 2;
            else if (p_case == 4 && dport == 0) fl |= 2;
            else if (p_case == 4 && sport == 0) fl |= 4;
            ev.word_port[i] = word_port;
            ev.flags[i] = fl;
            memcpy(w, "-1_", 3);
            const int wl = flow_word(ev, i, w + 3);
            const sv plain(w + 3, wl), marked(w, wl + 3);
            ev.a_doc[i] = lookup(ev.docs, f[8]);   // :228 row(8)
            ev.b_doc[i] = lookup(ev.docs, f[9]);   // :230 row(9)
            ev.a_word[i] = lookup(ev.words, (fl & 2) ? marked : plain);  // :229 row(33) src_word
            ev.b_word[i] = lookup(ev.words, (fl & 4) ? marked : plain);  // :231 row(34) dest_word
        }
    });
    laps.lap("bins + words + lookups");
    return rc;
}

// the featurised row as the script holds it just before scoring (columns 0..34)
void flow_row_text(const oni_events &ev, size_t i, std::string &out)
{
    char num[64], w[160];
    std::vector<sv> f;
    split_java(ev.row[i], ',', f);
    out.assign(ev.row[i]);
    out += ',';
    java_double_str(num, ev.num_time[i]);
    out += num;
    for (int b = 0; b < 3; b++) { out += ','; out += std::to_string((int)ev.bins[3 * i + b]); }
    out += ',';
    java_double_str(num, ev.word_port[i]);
    out += num;
    out += ',';
    const bool fwd = ev.flags[i] & 1;
    out += fwd ? f[8] : f[9];
    out += ' ';
    out += fwd ? f[9] : f[8];
    flow_word(ev, i, w);
    out += ',';
    if (ev.flags[i] & 2) out += "-1_";
    out += w;
    out += ',';
    if (ev.flags[i] & 4) out += "-1_";
    out += w;
}

// ------------------------------------------------------------------ dns
enum { C_TIME = 1, C_LEN = 2, C_IPDST = 3, C_NAME = 4, C_TYPE = 6, C_RCODE = 7, DNS_COLS = 8 };

// entropy (dns_post_lda.scala:206-212): sum over the distinct characters of -p*log10(p)/log10(2).
// The script adds the terms in the iteration order of a Scala hash map; here: ascending byte value.
double shannon(sv s)
{
    int cnt[256] = {0};
    for (unsigned char c : s) cnt[c]++;
    const double len = (double)s.size();
    double sum = 0.0;
    for (int c = 0; c < 256; c++)
        if (cnt[c]) {
            const double p = (double)cnt[c] / len;
            sum += -p * log10(p) / log10(2.0);
        }
    return sum;
}

int dns_word(const oni_events &ev, size_t i, const std::vector<sv> &f, char *out)
{
    const uint8_t *b = &ev.bins[5 * i];
    // :282-287 top_domain _ frame_len _ unix_tstamp _ subdomain.length _ entropy _ num.periods _ type _ rcode
    int n = sprintf(out, "%d_%d_%d_%d_%d_%d_", (int)ev.top[i], (int)b[1], (int)b[0], (int)b[2], (int)b[3], (int)b[4]);
    if (f[C_TYPE].size() + f[C_RCODE].size() > 100) return -1;
    memcpy(out + n, f[C_TYPE].data(), f[C_TYPE].size()); n += (int)f[C_TYPE].size();
    out[n++] = '_';
    memcpy(out + n, f[C_RCODE].data(), f[C_RCODE].size()); n += (int)f[C_RCODE].size();
    out[n] = 0;
    return n;
}

int dns_open(oni_events &ev)
{
    // :76-81 top-1m.csv: the first label of the second column
    {
        std::vector<sv> lines, f, lab;
        split_lines(ev.top_text, lines);
        for (size_t i = 0; i < lines.size(); i++) {
            split_java(lines[i], ',', f);
            if (f.size() < 2) return oni_set_error(ONI_ERR_IO, "top domains line %zu: no second column", i + 1);
            split_java(f[1], '.', lab);
            if (lab.empty()) return oni_set_error(ONI_ERR_IO, "top domains line %zu: empty domain", i + 1);
            ev.top_domains.insert(lab[0]);
        }
        for (const char *c : COUNTRY_CODES) ev.country.insert(sv(c));
    }
    std::vector<sv> lines;
    split_lines(ev.data, lines);
    ev.lines_in = (int64_t)lines.size();
    // :138 rows with exactly as many fields as the selected columns
    {
        std::vector<uint8_t> keep(lines.size(), 0);
        ONI_TRY(parallel_rows(lines.size(), [&](size_t lo, size_t hi, int, std::string &) {
            std::vector<sv> f;
            for (size_t i = lo; i < hi; i++) {
                split_java(lines[i], ',', f);
                keep[i] = f.size() == DNS_COLS;
            }
        }));
        size_t m = 0;
        for (size_t i = 0; i < lines.size(); i++)
            if (keep[i]) lines[m++] = lines[i];
        lines.resize(m);
    }
    ev.row.swap(lines);
    const size_t m = ev.row.size();
    ev.domain.resize(m); ev.subdomain.resize(m); ev.sub_len.resize(m); ev.num_parts.resize(m);
    ev.entropy.resize(m); ev.top.resize(m); ev.bins.resize(5 * m);
    std::vector<double> tstamp(m), flen(m);
    ONI_TRY(parallel_rows(m, [&](size_t lo, size_t hi, int, std::string &err) {
        std::vector<sv> f, part;
        for (size_t i = lo; i < hi; i++) {
            split_java(ev.row[i], ',', f);
            if (!parse_double(f[C_TIME], &tstamp[i]) || !parse_double(f[C_LEN], &flen[i])) {
                if (err.empty()) err = "dns row " + std::to_string(i + 1) + ": unix_tstamp / frame_len does not parse";
                return;
            }
            // extract_subdomain (:154-190)
            split_java(f[C_NAME], '.', part);
            const int np = (int)part.size();
            std::string domain = "None", sub = "None";
            const bool is_ip = np > 2 && part[np - 1] == "arpa" && part[np - 2] == "in-addr";
            if (np > 2 && !is_ip) {
                const int cut = ev.country.count(part[np - 1]) ? np - 3 : np - 2;
                domain.assign(part[cut]);
                if (cut >= 1) {
                    sub.clear();
                    for (int j = 0; j < cut; j++) { if (j) sub += '.'; sub += part[j]; }
                }
            }
            ev.sub_len[i] = (sub != "None") ? (int32_t)sub.size() : 0;
            ev.num_parts[i] = np;
            ev.entropy[i] = shannon(sub);  // :243 of the sub-domain text, "None" included
            // :272 top_domain
            ev.top[i] = domain == "intel" ? 2 : (ev.top_domains.count(sv(domain)) ? 1 : 0);
            ev.domain[i].swap(domain);
            ev.subdomain[i].swap(sub);
        }
    }));
    // :249-266 cuts; the last three over the positive values only
    ev.ncuts[0] = 10; ev.ncuts[1] = 10; ev.ncuts[2] = 5; ev.ncuts[3] = 5; ev.ncuts[4] = 5;
    {
        std::vector<double> tmp(tstamp);
        ecdf_cuts(tmp, QUANT10, 10, ev.cuts[0]);
        tmp = flen;
        ecdf_cuts(tmp, QUANT10, 10, ev.cuts[1]);
        tmp.clear();
        for (size_t i = 0; i < m; i++) if (ev.sub_len[i] > 0) tmp.push_back((double)ev.sub_len[i]);
        ecdf_cuts(tmp, QUINT5, 5, ev.cuts[2]);
        tmp.clear();
        for (size_t i = 0; i < m; i++) if (ev.entropy[i] > 0) tmp.push_back(ev.entropy[i]);
        ecdf_cuts(tmp, QUINT5, 5, ev.cuts[3]);
        tmp.clear();
        for (size_t i = 0; i < m; i++) if (ev.num_parts[i] > 0) tmp.push_back((double)ev.num_parts[i]);
        ecdf_cuts(tmp, QUINT5, 5, ev.cuts[4]);
    }
    ev.a_doc.resize(m); ev.a_word.resize(m);
    return parallel_rows(m, [&](size_t lo, size_t hi, int, std::string &err) {
        std::vector<sv> f;
        char w[256];
        for (size_t i = lo; i < hi; i++) {
            split_java(ev.row[i], ',', f);
            uint8_t *b = &ev.bins[5 * i];
            b[0] = (uint8_t)bin_of(tstamp[i], ev.cuts[0], 10);
            b[1] = (uint8_t)bin_of(flen[i], ev.cuts[1], 10);
            b[2] = (uint8_t)bin_of((double)ev.sub_len[i], ev.cuts[2], 5);
            b[3] = (uint8_t)bin_of(ev.entropy[i], ev.cuts[3], 5);
            b[4] = (uint8_t)bin_of((double)ev.num_parts[i], ev.cuts[4], 5);
            const int wl = dns_word(ev, i, f, w);
            if (wl < 0) { if (err.empty()) err = "dns row " + std::to_string(i + 1) + ": query type / rcode too long"; return; }
            ev.a_doc[i] = lookup(ev.docs, f[C_IPDST]);  // :313
            ev.a_word[i] = lookup(ev.words, sv(w, wl));   // :314
        }
    });
}

// columns 0..14 of the script's row just before scoring
void dns_row_text(const oni_events &ev, size_t i, std::string &out)
{
    char num[64], w[256];
    std::vector<sv> f;
    split_java(ev.row[i], ',', f);
    out.assign(ev.row[i]);
    out += ','; out += ev.domain[i];
    out += ','; out += ev.subdomain[i];
    out += ','; out += std::to_string(ev.sub_len[i]);
    out += ','; out += std::to_string(ev.num_parts[i]);
    java_double_str(num, ev.entropy[i]);
    out += ','; out += num;
    out += ','; out += std::to_string((int)ev.top[i]);
    dns_word(ev, i, f, w);
    out += ','; out += w;
}

// Analyst feedback of the pre-LDA scripts: rows of the scores file that an analyst marked
// severity 3 ("not a threat") are turned back into event rows and appended `dup` times each, so
// that the next model finds them ordinary (flow_pre_lda.scala:147-263, dns_pre_lda.scala:70-138).
bool parse_int_java(sv f, long *out)  // String.toInt: optional sign, digits, nothing else
{
    if (f.empty() || f.size() > 11) return false;
    char buf[16];
    memcpy(buf, f.data(), f.size());
    buf[f.size()] = 0;
    char *end = nullptr;
    const long v = strtol(buf, &end, 10);
    const bool starts = (buf[0] >= '0' && buf[0] <= '9') || ((buf[0] == '-' || buf[0] == '+') && buf[1] >= '0' && buf[1] <= '9');
    if (!starts || end == buf || *end) return false;
    *out = v;
    return true;
}

int append_feedback(int kind, const char *path, int dup, std::string &data)
{
    if (dup < 0) return oni_set_error(ONI_ERR_ARG, "feedback: duplication factor %d", dup);
    std::string text;
    ONI_TRY(slurp(path, text));
    std::vector<sv> lines, f, a, b;
    split_lines(text, lines);
    if (!data.empty() && data.back() != '\n') data += '\n';
    std::string row;
    for (size_t i = 1; i < lines.size(); i++) {  // (the first line is the header)
        split_java(lines[i], ',', f);
        long sev = 0;
        row.clear();
        if (kind) {
            // flow_scores.csv: 22 columns; 0 severity, 1 start time "YYYY-MM-DD HH:MM:SS", 2/3 IPs,
            // 4/5 ports, 8 packets, 9 bytes.  Every other column of the flow row becomes "##".
            if (f.size() != 22) continue;
            if (!parse_int_java(f[0], &sev)) return oni_set_error(ONI_ERR_IO, "%s line %zu: severity is not an integer", path, i + 1);
            if (sev != 3) continue;
            split_java(f[1], ' ', a);
            if (a.size() < 2) return oni_set_error(ONI_ERR_IO, "%s line %zu: start time has no clock part", path, i + 1);
            split_java(a[1], ':', b);
            if (b.size() < 3) return oni_set_error(ONI_ERR_IO, "%s line %zu: clock part is not HH:MM:SS", path, i + 1);
            for (int c = 0; c < 27; c++) {
                sv v = "##";
                if (c == 4) v = b[0]; else if (c == 5) v = b[1]; else if (c == 6) v = b[2];
                else if (c == 8) v = f[2]; else if (c == 9) v = f[3]; else if (c == 10) v = f[4];
                else if (c == 11) v = f[5]; else if (c == 16) v = f[8]; else if (c == 17) v = f[9];
                if (c) row += ',';
                row.append(v.data(), v.size());
            }
        } else {
            // dns_scores.csv: 0 frame_time, 1 frame_len, 2 ip_dst, 3 name, 4 class, 5 type, 6 rcode,
            // 18 severity, 23 unix_tstamp
            if (f.size() < 24) return oni_set_error(ONI_ERR_IO, "%s line %zu: %zu columns, 24 expected", path, i + 1, f.size());
            long flen = 0;
            if (!parse_int_java(trim_java(f[1]), &flen) || !parse_int_java(trim_java(f[18]), &sev))
                return oni_set_error(ONI_ERR_IO, "%s line %zu: frame_len / severity is not an integer", path, i + 1);
            if (sev != 3) continue;
            row.append(f[0].data(), f[0].size()); row += ',';
            row.append(f[23].data(), f[23].size()); row += ',';
            row += std::to_string(flen);
            for (int c = 2; c <= 6; c++) { row += ','; row.append(f[c].data(), f[c].size()); }
        }
        for (int d = 0; d < dup; d++) { data += row; data += '\n'; }
    }
    return ONI_OK;
}

int open_common(int kind, const char *events_csv, const char *top_csv, const char *doc_results,
                const char *word_results, oni_events **out, const char *feedback_csv = nullptr, int dup = 0)
{
    if (!events_csv || !out || (kind == 0 && !top_csv) || (!doc_results) != (!word_results))
        return oni_set_error(ONI_ERR_ARG, "events_open: null argument");
    *out = nullptr;
    oni_events *ev = new (std::nothrow) oni_events();
    if (!ev) return oni_set_error(ONI_ERR_NOMEM, "events_open: out of memory");
    ev->kind = kind;
    int rc = ONI_OK;
    Laps laps;
    try {
        rc = slurp(events_csv, ev->data);
        if (rc == ONI_OK && feedback_csv) rc = append_feedback(kind, feedback_csv, dup, ev->data);
        laps.lap("read events file");
        if (rc == ONI_OK && kind == 0) rc = slurp(top_csv, ev->top_text);
        ev->has_tables = doc_results != nullptr;  // absent for the pre-LDA word count
        if (rc == ONI_OK && ev->has_tables) rc = load_table(doc_results, ev->docs);
        if (rc == ONI_OK && ev->has_tables) rc = load_table(word_results, ev->words);
        if (rc == ONI_OK) rc = kind ? flow_open(*ev) : dns_open(*ev);
    } catch (const std::bad_alloc &) {
        rc = oni_set_error(ONI_ERR_NOMEM, "events_open: out of memory");
    }
    if (rc != ONI_OK) { delete ev; return rc; }
    *out = ev;
    return ONI_OK;
}

}  // namespace

// ------------------------------------------------------------------ C ABI
extern "C" int oni_flow_events_open(const char *flow_csv, const char *doc_results_csv,
                                    const char *word_results_csv, oni_events **out)
{
    return open_common(1, flow_csv, nullptr, doc_results_csv, word_results_csv, out);
}

extern "C" int oni_dns_events_open(const char *dns_csv, const char *top_domains_csv, const char *doc_results_csv,
                                   const char *word_results_csv, oni_events **out)
{
    return open_common(0, dns_csv, top_domains_csv, doc_results_csv, word_results_csv, out);
}

extern "C" int oni_flow_events_open_feedback(const char *flow_csv, const char *feedback_csv, int32_t dup_factor,
                                             const char *doc_results_csv, const char *word_results_csv,
                                             oni_events **out)
{
    if (!feedback_csv) return oni_set_error(ONI_ERR_ARG, "events_open_feedback: null feedback file");
    return open_common(1, flow_csv, nullptr, doc_results_csv, word_results_csv, out, feedback_csv, dup_factor);
}

extern "C" int oni_dns_events_open_feedback(const char *dns_csv, const char *top_domains_csv, const char *feedback_csv,
                                            int32_t dup_factor, const char *doc_results_csv,
                                            const char *word_results_csv, oni_events **out)
{
    if (!feedback_csv) return oni_set_error(ONI_ERR_ARG, "events_open_feedback: null feedback file");
    return open_common(0, dns_csv, top_domains_csv, doc_results_csv, word_results_csv, out, feedback_csv, dup_factor);
}

extern "C" void oni_events_close(oni_events *ev) { delete ev; }

extern "C" int oni_events_info(const oni_events *ev, int64_t *lines_in, int64_t *rows, int32_t *num_docs,
                               int32_t *num_words)
{
    if (!ev) return oni_set_error(ONI_ERR_ARG, "events_info: null handle");
    if (lines_in) *lines_in = ev->lines_in;
    if (rows) *rows = (int64_t)ev->row.size();
    if (num_docs) *num_docs = ev->docs.rows;
    if (num_words) *num_words = ev->words.rows;
    return ONI_OK;
}

extern "C" int oni_events_cuts(const oni_events *ev, int32_t which, double *out, int32_t cap, int32_t *n)
{
    if (!ev || !out || which < 0 || which >= 5 || ev->ncuts[which] == 0 || cap < ev->ncuts[which])
        return oni_set_error(ONI_ERR_ARG, "events_cuts: bad cut vector %d or capacity %d", which, cap);
    memcpy(out, ev->cuts[which], sizeof(double) * ev->ncuts[which]);
    if (n) *n = ev->ncuts[which];
    return ONI_OK;
}

extern "C" int oni_events_indices(const oni_events *ev, int32_t *a_doc, int32_t *b_doc, int32_t *a_word,
                                  int32_t *b_word)
{
    if (!ev) return oni_set_error(ONI_ERR_ARG, "events_indices: null handle");
    const size_t bytes = ev->row.size() * sizeof(int32_t);
    if (a_doc) memcpy(a_doc, ev->a_doc.data(), bytes);
    if (a_word) memcpy(a_word, ev->a_word.data(), bytes);
    if (ev->kind) {
        if (b_doc) memcpy(b_doc, ev->b_doc.data(), bytes);
        if (b_word) memcpy(b_word, ev->b_word.data(), bytes);
    }
    return ONI_OK;
}

extern "C" int64_t oni_events_row_text(const oni_events *ev, int64_t row, char *buf, int64_t cap)
{
    if (!ev || row < 0 || row >= (int64_t)ev->row.size()) {
        oni_set_error(ONI_ERR_ARG, "events_row_text: row out of range");
        return -1;
    }
    std::string s;
    if (ev->kind) flow_row_text(*ev, (size_t)row, s); else dns_row_text(*ev, (size_t)row, s);
    if (buf && cap > 0) {
        const size_t c = std::min<size_t>(s.size(), (size_t)cap - 1);
        memcpy(buf, s.data(), c);
        buf[c] = 0;
    }
    return (int64_t)s.size();
}

// filter + rank on the GPU, then the text of the surviving rows in rank order
// (flow_post_lda.scala:238-248, dns_post_lda.scala:319-330)
extern "C" int oni_events_score_write(oni_events *ev, double tol, const char *out_path, int32_t device,
                                      int64_t *num_scored)
{
    if (!ev || !out_path) return oni_set_error(ONI_ERR_ARG, "events_score_write: null argument");
    if (!ev->has_tables)
        return oni_set_error(ONI_ERR_ARG, "events_score_write: opened without doc_results.csv / word_results.csv");
    const int64_t n = (int64_t)ev->row.size();
    if (n == 0) {  // nothing survived the filters: an empty result, no GPU work
        FILE *f0 = fopen(out_path, "w");
        if (!f0) return oni_set_error(ONI_ERR_IO, "cannot write %s: %s", out_path, strerror(errno));
        fclose(f0);
        if (num_scored) *num_scored = 0;
        return ONI_OK;
    }
    std::vector<int64_t> idx((size_t)std::max<int64_t>(n, 1));
    std::vector<double> key(idx.size()), sa(idx.size()), sb(idx.size());
    int64_t total = 0;
    if (ev->kind)
        ONI_TRY(oni_score_flow(device, 20, ev->docs.rows, ev->words.rows, ev->docs.values.data(),
                               ev->words.values.data(), n, ev->a_doc.data(), ev->b_doc.data(), ev->a_word.data(),
                               ev->b_word.data(), tol, n, idx.data(), key.data(), sa.data(), sb.data(), &total));
    else
        ONI_TRY(oni_score_dns(device, 20, ev->docs.rows, ev->words.rows, ev->docs.values.data(),
                              ev->words.values.data(), n, ev->a_doc.data(), ev->a_word.data(), tol, n, idx.data(),
                              key.data(), &total));
    FILE *f = fopen(out_path, "w");
    if (!f) return oni_set_error(ONI_ERR_IO, "cannot write %s: %s", out_path, strerror(errno));
    setvbuf(f, nullptr, _IOFBF, 1 << 20);
    std::string s;
    char num[64];
    for (int64_t r = 0; r < total; r++) {
        const size_t i = (size_t)idx[r];
        if (ev->kind) {
            flow_row_text(*ev, i, s);
            java_double_str(num, sa[r]); s += ','; s += num;
            java_double_str(num, sb[r]); s += ','; s += num;
        } else {
            dns_row_text(*ev, i, s);
            java_double_str(num, key[r]); s += ','; s += num;
        }
        s += '\n';
        fwrite(s.data(), 1, s.size(), f);
    }
    if (fclose(f) != 0) return oni_set_error(ONI_ERR_IO, "write failed on %s", out_path);
    if (num_scored) *num_scored = total;
    return ONI_OK;
}

// The word-count tail of flow_pre_lda.scala:371-382 and dns_pre_lda.scala:326-330: one line
// "ip,word,count" per distinct (ip, word).  Flow counts the source side (row(8), src_word) and the
// destination side (row(9), dest_word) separately and writes both lists, so a pair can appear
// twice (lda_pre.py keeps both).  Spark leaves the line order to its hash partitioner; here:
// first appearance, source-side list first.
namespace {
// (ip, word) without building the word's text: the word is a function of these fields
struct WKey {
    sv ip, type, rcode;  // type / rcode: dns only
    uint64_t u;          // flow: bits of word_port
    uint32_t v;          // bins, flags
    bool operator==(const WKey &o) const
    {
        return u == o.u && v == o.v && ip == o.ip && type == o.type && rcode == o.rcode;
    }
};
struct WHash {
    size_t operator()(const WKey &k) const
    {
        const std::hash<sv> h;
        uint64_t x = h(k.ip) * 0x9E3779B97F4A7C15ull ^ k.u;
        x = (x ^ (x >> 29)) * 0xBF58476D1CE4E5B9ull ^ k.v;
        if (!k.type.empty() || !k.rcode.empty()) x = (x ^ h(k.type)) * 0x94D049BB133111EBull ^ h(k.rcode);
        return (size_t)(x ^ (x >> 32));
    }
};
struct WEntry { WKey key; size_t first_row; int64_t count; };
struct WTable {
    std::unordered_map<WKey, size_t, WHash> slot;
    std::vector<WEntry> order;  // first appearance
    void add(const WKey &k, size_t row, int64_t n)
    {
        auto it = slot.find(k);
        if (it == slot.end()) { slot.emplace(k, order.size()); order.push_back(WEntry{k, row, n}); }
        else order[it->second].count += n;
    }
};
}  // namespace

extern "C" int oni_events_word_counts(const oni_events *ev, const char *out_path, int64_t *num_lines)
{
    if (!ev || !out_path) return oni_set_error(ONI_ERR_ARG, "events_word_counts: null argument");
    FILE *f = fopen(out_path, "w");
    if (!f) return oni_set_error(ONI_ERR_IO, "cannot write %s: %s", out_path, strerror(errno));
    setvbuf(f, nullptr, _IOFBF, 1 << 20);
    const size_t n = ev->row.size();
    int64_t lines = 0;
    Laps laps;
    for (int side = 0; side < (ev->kind ? 2 : 1); side++) {
        // keys of all rows side by side, then every thread counts the keys whose hash falls in its
        // share; sorting the distinct pairs by the row that introduced them gives first-appearance order
        std::vector<WKey> keys(n);
        std::vector<uint32_t> hv(n);
        ONI_TRY(parallel_rows(n, [&](size_t lo, size_t hi, int, std::string &) {
            std::vector<sv> fld;
            const WHash hasher;
            for (size_t i = lo; i < hi; i++) {
                split_java(ev->row[i], ',', fld);
                WKey k{};
                if (ev->kind) {
                    k.ip = fld[8 + side];
                    memcpy(&k.u, &ev->word_port[i], 8);
                    k.v = (uint32_t)ev->bins[3 * i] | (uint32_t)ev->bins[3 * i + 1] << 8 |
                          (uint32_t)ev->bins[3 * i + 2] << 16 | (uint32_t)((ev->flags[i] >> (side ? 2 : 1)) & 1) << 24;
                } else {
                    const uint8_t *b = &ev->bins[5 * i];
                    k.ip = fld[C_IPDST]; k.type = fld[C_TYPE]; k.rcode = fld[C_RCODE];
                    k.u = (uint64_t)b[0] | (uint64_t)b[1] << 8 | (uint64_t)b[2] << 16 | (uint64_t)b[3] << 24 |
                          (uint64_t)b[4] << 32;
                    k.v = ev->top[i];
                }
                keys[i] = k;
                const size_t h = hasher(k);
                hv[i] = (uint32_t)(h ^ (h >> 32));
            }
        }));
        laps.lap("  row keys");
        const int T = io_threads();
        std::vector<std::vector<WEntry>> found(T);
        {
            std::vector<std::thread> th;
            for (int t = 0; t < T; t++)
                th.emplace_back([&, t] {
                    WTable mine;
                    for (size_t i = 0; i < n; i++)
                        if ((int)((hv[i] >> 8) % (uint32_t)T) == t) mine.add(keys[i], i, 1);
                    found[t].swap(mine.order);
                });  // (the table is torn down by its own thread)
            for (auto &x : th) x.join();
        }
        struct { std::vector<WEntry> order; } all;
        for (int t = 0; t < T; t++) all.order.insert(all.order.end(), found[t].begin(), found[t].end());
        std::sort(all.order.begin(), all.order.end(),
                  [](const WEntry &x, const WEntry &y) { return x.first_row < y.first_row; });
        laps.lap("  count + order");
        std::vector<sv> fld;
        char w[256];
        for (const WEntry &e : all.order) {
            const size_t i = e.first_row;
            fwrite(e.key.ip.data(), 1, e.key.ip.size(), f);
            fputc(',', f);
            if (ev->kind) {
                if (ev->flags[i] & (side ? 4 : 2)) fputs("-1_", f);
                flow_word(*ev, i, w);
            } else {
                split_java(ev->row[i], ',', fld);
                dns_word(*ev, i, fld, w);
            }
            fprintf(f, "%s,%lld\n", w, (long long)e.count);
        }
        lines += (int64_t)all.order.size();
        laps.lap("  write");
    }
    if (fclose(f) != 0) return oni_set_error(ONI_ERR_IO, "write failed on %s", out_path);
    laps.lap("word counts");
    if (num_lines) *num_lines = lines;
    return ONI_OK;
}

// test hook: Double.toString restatement
extern "C" int oni_internal_java_double_str(double v, char *out) { return java_double_str(out, v); }
