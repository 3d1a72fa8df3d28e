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

}  // namespace

extern "C" int oni_lda_main(int argc, char **argv)
{
    if (argc > 1 && strcmp(argv[1], "est") == 0) return run_est(argc, argv);
    if (argc > 1 && strcmp(argv[1], "inf") == 0) return run_inf(argc, argv);
    // lda-estimate.c:464-465
    printf("usage : lda est [alpha] [k] [settings] [#gpus] [data] [random/seeded/*] [output directory]\n");
    printf("        lda inf [settings] [model] [data] [output filename]\n");
    return ONI_OK;
}
