// lda_cli.cu -- the reference's process contract on top of the engine (host code).
//
// oni_lda_main() accepts exactly the argv of oni-lda-c's main
// (lda-estimate.c:403-468) and produces the same files:
//
//   lda est <alpha> <k> <settings> <ngpus> <model.dat> <random|seeded|model_root> <outdir>
//       -> <outdir>/000.beta 000.other likelihood.dat final.beta final.other final.gamma
//   lda inf <settings> <model_root> <model.dat> <out_name>
//       -> <out_name>-lda-lhood.dat <out_name>-gamma.dat      (lda-estimate.c:358-388)
//
// ml_ops.sh:80 passes PROCESS_COUNT as argv[5]; here that number is the count of
// GPUs to shard the documents over (clamped to the visible devices).
#include <errno.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <time.h>

#include <string>
#include <vector>

#include "common.h"
#include "mt19937.h"

namespace {

// load_lda_model (lda-model.c:220-251): alpha and every beta value pass through
// a float, exactly as the reference's fscanf("%f") does.
int load_model(const char *root, int *K, int *V, double *alpha, std::vector<double> &logbeta)
{
    std::string name = std::string(root) + ".other";
    FILE *f = fopen(name.c_str(), "r");
    if (!f) return oni_set_error(ONI_ERR_IO, "cannot open %s: %s", name.c_str(), strerror(errno));
    float a = 0;
    int ok = fscanf(f, "num_topics %d\n", K);
    ok += fscanf(f, "num_terms %d\n", V);
    ok += fscanf(f, "alpha %f\n", &a);
    fclose(f);
    if (ok != 3 || *K < 1 || *V < 1) return oni_set_error(ONI_ERR_IO, "%s: malformed", name.c_str());
    *alpha = a;
    name = std::string(root) + ".beta";
    f = fopen(name.c_str(), "r");
    if (!f) return oni_set_error(ONI_ERR_IO, "cannot open %s: %s", name.c_str(), strerror(errno));
    logbeta.resize((size_t)*K * *V);
    for (size_t i = 0; i < logbeta.size(); i++) {
        float x;
        if (fscanf(f, "%f", &x) != 1) {
            fclose(f);
            return oni_set_error(ONI_ERR_IO, "%s: expected %zu values, got %zu", name.c_str(), logbeta.size(), i);
        }
        logbeta[i] = x;
    }
    fclose(f);
    return ONI_OK;
}

// corpus_initialize_ss + lda_mle(model, ss, 0) (lda-model.c:120-144, 27-44)
void seeded_model(const oni_lda_corpus *c, int K, uint32_t seed, std::vector<double> &logbeta)
{
    const int V = c->num_terms;
    oni::CokusMT mt(seed);
    logbeta.assign((size_t)K * V, 0.0);
    std::vector<double> cw((size_t)V);
    for (int k = 0; k < K; k++) {
        std::fill(cw.begin(), cw.end(), 0.0);
        // NUM_INIT = 1 (lda-model.h:12)
        const int d = (int)floor(mt.uniform() * c->num_docs);
        printf("initialized with document %d\n", d);
        for (int64_t i = c->doc_ptr[d]; i < c->doc_ptr[d + 1]; i++) cw[c->word[i]] += c->count[i];
        double ct = 0.0;
        for (int n = 0; n < V; n++) {
            cw[n] += 1.0;
            ct = ct + cw[n];
        }
        const double lct = log(ct);
        for (int n = 0; n < V; n++) logbeta[(size_t)k * V + n] = (cw[n] > 0) ? log(cw[n]) - lct : -100.0;
    }
}

uint32_t pick_seed()
{
    const char *e = getenv("ONI_LDA_SEED");
    if (e && *e) return (uint32_t)strtoul(e, nullptr, 10);
    return (uint32_t)time(nullptr);  // lda-estimate.c:414-416
}

struct IterLog {
    FILE *f;
};
void on_iter(void *user, int iter, double lik, double conv)
{
    IterLog *log = (IterLog *)user;
    printf("**** em iteration %d ****  likelihood %10.10f\n", iter, lik);
    if (log->f) {
        fprintf(log->f, "%10.10f\t%5.5e\n", lik, conv);  // lda-estimate.c:266
        fflush(log->f);
    }
}

int run_est(int argc, char **argv)
{
    if (argc < 9) return oni_set_error(ONI_ERR_ARG, "est: expected 7 arguments");
    const float alpha_f = (float)atof(argv[2]);  // lda-estimate.c:425: parsed into a float
    const int K = atoi(argv[3]);
    const char *settings_path = argv[4];
    int ngpus = atoi(argv[5]);
    const char *corpus_path = argv[6];
    const char *start = argv[7];
    const char *outdir = argv[8];

    oni_lda_settings st;
    ONI_TRY(oni_lda_read_settings(settings_path, &st));
    printf("reading data from %s\n", corpus_path);
    oni_lda_corpus corpus;
    ONI_TRY(oni_lda_read_corpus(corpus_path, &corpus));
    printf("number of docs    : %d\n", corpus.num_docs);
    printf("number of terms   : %d\n", corpus.num_terms);
    mkdir(outdir, S_IRUSR | S_IWUSR | S_IXUSR);  // utils.c:86-89

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        oni_lda_free_corpus(&corpus);
        return oni_set_error(ONI_ERR_CUDA, "no CUDA device: this engine has no CPU path");
    }
    if (ngpus < 1) ngpus = 1;
    if (ngpus > ndev) {
        printf("note: %d workers requested, %d GPUs visible; using %d\n", ngpus, ndev, ndev);
        ngpus = ndev;
    }

    int Kmodel = K, V = corpus.num_terms;
    double alpha = alpha_f;
    std::vector<double> lb0;
    const uint32_t seed = pick_seed();
    int rc = ONI_OK;
    if (strcmp(start, "seeded") == 0) {
        seeded_model(&corpus, K, seed, lb0);
    } else if (strcmp(start, "random") == 0) {
        oni_internal_random_model(K, V, seed, lb0);
    } else {
        printf("loading %s\n", start);
        rc = load_model(start, &Kmodel, &V, &alpha, lb0);  // alpha comes from the file (lda-estimate.c:163-167)
        if (rc == ONI_OK && V < corpus.num_terms)
            rc = oni_set_error(ONI_ERR_ARG, "model has %d terms, corpus needs %d", V, corpus.num_terms);
        if (rc == ONI_OK) corpus.num_terms = V;
    }
    if (rc != ONI_OK) { oni_lda_free_corpus(&corpus); return rc; }

    std::string root = std::string(outdir) + "/000";
    rc = oni_lda_save_model(root.c_str(), Kmodel, V, lb0.data(), alpha);  // lda-estimate.c:169-170
    IterLog log{nullptr};
    if (rc == ONI_OK) {
        std::string lname = std::string(outdir) + "/likelihood.dat";
        log.f = fopen(lname.c_str(), "w");
        if (!log.f) rc = oni_set_error(ONI_ERR_IO, "cannot write %s: %s", lname.c_str(), strerror(errno));
    }
    std::vector<double> lb((size_t)Kmodel * V), gamma((size_t)corpus.num_docs * Kmodel);
    oni_lda_result res;
    memset(&res, 0, sizeof res);
    res.logbeta = lb.data();
    res.gamma = gamma.data();
    if (rc == ONI_OK)
        rc = oni_internal_estimate(&corpus, Kmodel, alpha, &st, lb0.data(), ngpus, &res, on_iter, &log);
    if (log.f) fclose(log.f);
    if (rc == ONI_OK) {
        root = std::string(outdir) + "/final";
        rc = oni_lda_save_model(root.c_str(), Kmodel, V, lb.data(), res.alpha);  // :274-275
    }
    if (rc == ONI_OK) {
        printf("*****saving gamma*****\n");
        std::string gname = std::string(outdir) + "/final.gamma";
        rc = oni_lda_save_gamma(gname.c_str(), corpus.num_docs, Kmodel, gamma.data());  // :313-317
    }
    if (rc == ONI_OK)
        printf("em: %d iterations in %.3f s on %d GPU(s)\n", res.em_iters, res.seconds_em, ngpus);
    oni_lda_free_corpus(&corpus);
    return rc;
}

int run_inf(int argc, char **argv)
{
    if (argc < 6) return oni_set_error(ONI_ERR_ARG, "inf: expected 4 arguments");
    oni_lda_settings st;
    ONI_TRY(oni_lda_read_settings(argv[2], &st));
    int K = 0, V = 0;
    double alpha = 0;
    std::vector<double> lb;
    printf("loading %s\n", argv[3]);
    ONI_TRY(load_model(argv[3], &K, &V, &alpha, lb));
    oni_lda_corpus corpus;
    ONI_TRY(oni_lda_read_corpus(argv[4], &corpus));
    int rc = ONI_OK;
    if (corpus.num_terms > V) rc = oni_set_error(ONI_ERR_ARG, "corpus has word ids beyond the model's %d terms", V);
    oni_lda_ctx *ctx = nullptr;
    if (rc == ONI_OK) rc = oni_lda_create(0, K, V, &st, &ctx);
    if (rc == ONI_OK) rc = oni_lda_set_corpus(ctx, corpus.num_docs, corpus.nnz, corpus.doc_ptr, corpus.word, corpus.count);
    if (rc == ONI_OK) rc = oni_lda_set_model(ctx, lb.data(), alpha);
    if (rc == ONI_OK) rc = oni_lda_e_step(ctx);
    std::vector<double> lik((size_t)corpus.num_docs), gamma((size_t)corpus.num_docs * K);
    if (rc == ONI_OK) rc = oni_lda_get_doc_likelihood(ctx, lik.data());
    if (rc == ONI_OK) rc = oni_lda_get_gamma(ctx, gamma.data());
    // assert(!isnan(likelihood)), lda-inference.c:78: the reference aborts, this returns ONI_ERR_NAN
    for (int d = 0; rc == ONI_OK && d < corpus.num_docs; d++)
        if (lik[d] != lik[d]) rc = oni_set_error(ONI_ERR_NAN, "the bound of document %d is NaN", d);
    if (rc == ONI_OK) {
        std::string name = std::string(argv[5]) + "-lda-lhood.dat";  // lda-estimate.c:374-375
        FILE *f = fopen(name.c_str(), "w");
        if (!f) rc = oni_set_error(ONI_ERR_IO, "cannot write %s: %s", name.c_str(), strerror(errno));
        else {
            for (int d = 0; d < corpus.num_docs; d++) fprintf(f, "%5.5f\n", lik[d]);  // :387
            fclose(f);
        }
    }
    if (rc == ONI_OK) {
        std::string name = std::string(argv[5]) + "-gamma.dat";  // :390-391
        rc = oni_lda_save_gamma(name.c_str(), corpus.num_docs, K, gamma.data());
    }
    oni_lda_destroy(ctx);
    oni_lda_free_corpus(&corpus);
    return rc;
}

// Python-2 str(float) / old-numpy str(float64): 12 significant digits, and ".0" appended when
// the text would otherwise look like an integer (lda_post.py:50,119-121)
int py2_float_str(char *out, double v)
{
    int n = snprintf(out, 48, "%.12g", v);
    bool plain = true;
    for (int i = 0; i < n; i++)
        if (out[i] == '.' || out[i] == 'e' || out[i] == 'n' || out[i] == 'i') { plain = false; break; }
    if (plain) { out[n++] = '.'; out[n++] = '0'; out[n] = 0; }
    return n;
}

// whitespace-separated doubles of a text matrix; *cols from the first line
int read_text_matrix(const char *path, std::vector<double> &a, int *rows, int *cols)
{
    FILE *f = fopen(path, "r");
    if (!f) return oni_set_error(ONI_ERR_IO, "cannot open %s: %s", path, strerror(errno));
    fseek(f, 0, SEEK_END);
    const long len = ftell(f);
    rewind(f);
    std::string buf((size_t)(len > 0 ? len : 0), '\0');
    const size_t got = fread(&buf[0], 1, buf.size(), f);
    fclose(f);
    buf.resize(got);
    *rows = 0; *cols = 0;
    const char *p = buf.c_str();
    int c = 0;
    while (*p) {
        while (*p == ' ' || *p == '\t' || *p == '\r') p++;
        if (*p == '\n') { if (c) { if (!*cols) *cols = c; else if (c != *cols) return oni_set_error(ONI_ERR_IO, "%s: ragged rows", path); (*rows)++; c = 0; } p++; continue; }
        if (!*p) break;
        char *e;
        const double v = strtod(p, &e);
        if (e == p) return oni_set_error(ONI_ERR_IO, "%s: not a number near byte %zu", path, (size_t)(p - buf.c_str()));
        a.push_back(v); c++; p = e;
    }
    if (c) { if (!*cols) *cols = c; else if (c != *cols) return oni_set_error(ONI_ERR_IO, "%s: ragged rows", path); (*rows)++; }
    return ONI_OK;
}

int read_lines(const char *path, std::vector<std::string> &lines)
{
    FILE *f = fopen(path, "r");
    if (!f) return oni_set_error(ONI_ERR_IO, "cannot open %s: %s", path, strerror(errno));
    char *line = nullptr; size_t cap = 0; ssize_t n;
    while ((n = getline(&line, &cap, f)) >= 0) {
        while (n > 0 && (line[n - 1] == '\n' || line[n - 1] == '\r')) n--;
        lines.emplace_back(line, (size_t)n);
    }
    free(line);
    fclose(f);
    return ONI_OK;
}

}  // namespace

// lda_post.py (ml_ops.sh:84): final.gamma / final.beta / doc.dat / words.dat under the prefix
// `dir` -> doc_results.csv ("ip,t0 t1 ...") and word_results.csv ("word,p0 p1 ...").
extern "C" int oni_lda_post_files(const char *dir, int32_t device)
{
    if (!dir) return oni_set_error(ONI_ERR_ARG, "lda_post_files: null directory prefix");
    const std::string d(dir);
    std::vector<double> gamma, beta;
    int D = 0, K = 0, Kb = 0, V = 0;
    ONI_TRY(read_text_matrix((d + "final.gamma").c_str(), gamma, &D, &K));
    ONI_TRY(read_text_matrix((d + "final.beta").c_str(), beta, &Kb, &V));
    if (K != Kb) return oni_set_error(ONI_ERR_IO, "final.gamma has %d topics, final.beta %d", K, Kb);
    std::vector<std::string> docs, words;
    ONI_TRY(read_lines((d + "doc.dat").c_str(), docs));
    ONI_TRY(read_lines((d + "words.dat").c_str(), words));
    if ((int)docs.size() < D) return oni_set_error(ONI_ERR_IO, "doc.dat has %zu lines, final.gamma %d rows", docs.size(), D);
    if ((int)words.size() < V) return oni_set_error(ONI_ERR_IO, "words.dat has %zu lines, final.beta %d columns", words.size(), V);
    std::vector<double> theta((size_t)D * K), pwz((size_t)V * K);
    ONI_TRY(oni_lda_post(device, D, K, V, gamma.data(), beta.data(), 0, theta.data(), pwz.data()));

    char num[64];
    FILE *f = fopen((d + "doc_results.csv").c_str(), "w");
    if (!f) return oni_set_error(ONI_ERR_IO, "cannot write %sdoc_results.csv: %s", dir, strerror(errno));
    setvbuf(f, nullptr, _IOFBF, 1 << 20);
    for (int i = 0; i < D; i++) {
        // doc = linecache.getline(doc_file, j).split(',')[1]   (lda_post.py:56)
        const std::string &ln = docs[i];
        const size_t c1 = ln.find(',');
        if (c1 == std::string::npos) { fclose(f); return oni_set_error(ONI_ERR_IO, "doc.dat line %d has no comma", i + 1); }
        const size_t c2 = ln.find(',', c1 + 1);
        fwrite(ln.data() + c1 + 1, 1, (c2 == std::string::npos ? ln.size() : c2) - c1 - 1, f);
        fputc(',', f);
        double total = 0;
        for (int k = 0; k < K; k++) total = total + gamma[(size_t)i * K + k];
        if (total > 0) {
            for (int k = 0; k < K; k++) {
                py2_float_str(num, theta[(size_t)i * K + k]);
                if (k) fputc(' ', f);
                fputs(num, f);
            }
        } else {
            // the script's literal fallback: twenty "0.0", whatever K is (lda_post.py:54)
            for (int k = 0; k < 20; k++) fputs(k ? " 0.0" : "0.0", f);
        }
        fputc('\n', f);
    }
    if (fclose(f) != 0) return oni_set_error(ONI_ERR_IO, "write failed on %sdoc_results.csv", dir);

    f = fopen((d + "word_results.csv").c_str(), "w");
    if (!f) return oni_set_error(ONI_ERR_IO, "cannot write %sword_results.csv: %s", dir, strerror(errno));
    setvbuf(f, nullptr, _IOFBF, 1 << 20);
    for (int w = 0; w < V; w++) {
        // word = '_'.join(fields[1:]) cut to 20 bytes (the "S20" array, lda_post.py:78-83)
        const std::string &ln = words[w];
        const size_t c1 = ln.find(',');
        std::string name = (c1 == std::string::npos) ? std::string() : ln.substr(c1 + 1);
        for (char &ch : name) if (ch == ',') ch = '_';
        if (name.size() > 20) name.resize(20);
        fwrite(name.data(), 1, name.size(), f);
        fputc(',', f);
        for (int k = 0; k < K; k++) {
            py2_float_str(num, pwz[(size_t)w * K + k]);
            if (k) fputc(' ', f);
            fputs(num, f);
        }
        fputc('\n', f);
    }
    if (fclose(f) != 0) return oni_set_error(ONI_ERR_IO, "write failed on %sword_results.csv", dir);
    return ONI_OK;
}

// <type>_post_lda.scala: HPATH holds doc_results.csv and word_results.csv; TOL is parsed like toDouble
static int run_events(int flow, const char *events, const char *top, const char *hpath, const char *tol,
                      const char *out_path)
{
    char *end = nullptr;
    const double t = strtod(tol, &end);
    if (end == tol || *end) return oni_set_error(ONI_ERR_ARG, "threshold '%s' is not a number", tol);
    const std::string h(hpath);
    const std::string docs = h + "doc_results.csv", words = h + "word_results.csv";
    oni_events *ev = nullptr;
    ONI_TRY(flow ? oni_flow_events_open(events, docs.c_str(), words.c_str(), &ev)
                 : oni_dns_events_open(events, top, docs.c_str(), words.c_str(), &ev));
    int64_t rows = 0, scored = 0;
    oni_events_info(ev, nullptr, &rows, nullptr, nullptr);
    const int rc = oni_events_score_write(ev, t, out_path, 0, &scored);
    oni_events_close(ev);
    if (rc == ONI_OK) printf("%lld of %lld events below %s -> %s\n", (long long)scored, (long long)rows, tol, out_path);
    return rc;
}

static int run_word_counts(int flow, const char *events, const char *top, const char *out_path,
                           const char *feedback, int dup)
{
    oni_events *ev = nullptr;
    if (feedback)  // LPATH/<type>_scores.csv and DUPFACTOR of the scripts
        ONI_TRY(flow ? oni_flow_events_open_feedback(events, feedback, dup, nullptr, nullptr, &ev)
                     : oni_dns_events_open_feedback(events, top, feedback, dup, nullptr, nullptr, &ev));
    else
        ONI_TRY(flow ? oni_flow_events_open(events, nullptr, nullptr, &ev)
                     : oni_dns_events_open(events, top, nullptr, nullptr, &ev));
    int64_t rows = 0, lines = 0;
    oni_events_info(ev, nullptr, &rows, nullptr, nullptr);
    const int rc = oni_events_word_counts(ev, out_path, &lines);
    oni_events_close(ev);
    if (rc == ONI_OK) printf("%lld events -> %lld (ip, word) pairs -> %s\n", (long long)rows, (long long)lines, out_path);
    return rc;
}

extern "C" int oni_lda_main(int argc, char **argv)
{
    if (argc > 1 && strcmp(argv[1], "est") == 0) return run_est(argc, argv);
    if (argc > 1 && strcmp(argv[1], "inf") == 0) return run_inf(argc, argv);
    // two additions to the reference's command line: the Python stages either side of the engine
    if (argc > 2 && strcmp(argv[1], "pre") == 0) return oni_lda_pre(argv[2]);            // lda_pre.py <dir>/
    if (argc > 2 && strcmp(argv[1], "post") == 0) return oni_lda_post_files(argv[2], 0);  // lda_post.py <dir>/
    // the Spark scoring stage (ml_ops.sh:108): <type>_post_lda.scala with FLOW_PATH / DNS_PATH, HPATH and TOL
    // the Spark word-count stage before lda_pre.py (ml_ops.sh:44-57): <type>_pre_lda.scala
    if (argc > 3 && strcmp(argv[1], "flow_pre") == 0)
        return run_word_counts(1, argv[2], nullptr, argv[3], argc > 5 ? argv[4] : nullptr, argc > 5 ? atoi(argv[5]) : 0);
    if (argc > 4 && strcmp(argv[1], "dns_pre") == 0)
        return run_word_counts(0, argv[2], argv[3], argv[4], argc > 6 ? argv[5] : nullptr, argc > 6 ? atoi(argv[6]) : 0);
    if (argc > 5 && strcmp(argv[1], "flow_post") == 0) return run_events(1, argv[2], nullptr, argv[3], argv[4], argv[5]);
    if (argc > 6 && strcmp(argv[1], "dns_post") == 0) return run_events(0, argv[2], argv[3], argv[4], argv[5], argv[6]);
    // lda-estimate.c:464-465
    printf("usage : lda est [alpha] [k] [settings] [#gpus] [data] [random/seeded/*] [output directory]\n");
    printf("        lda inf [settings] [model] [data] [output filename]\n");
    printf("        lda pre [dir/]    (lda_pre.py: doc_wc.dat -> words.dat doc.dat model.dat)\n");
    printf("        lda post [dir/]   (lda_post.py: final.gamma final.beta -> doc_results.csv word_results.csv)\n");
    printf("        lda flow_pre [flow.csv] [doc_wc.dat] ([flow_scores.csv] [dupfactor])             (flow_pre_lda.scala)\n");
    printf("        lda dns_pre [dns.csv] [top-1m.csv] [doc_wc.dat] ([dns_scores.csv] [dupfactor])   (dns_pre_lda.scala)\n");
    printf("        lda flow_post [flow.csv] [hpath/] [tol] [scored file]              (flow_post_lda.scala)\n");
    printf("        lda dns_post [dns.csv] [top-1m.csv] [hpath/] [tol] [scored file]   (dns_post_lda.scala)\n");
    return ONI_OK;
}
