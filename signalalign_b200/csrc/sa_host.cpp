// sa_host.cpp — host side of libsagpu: model files, integer preparation (band, split, kmer paths), small C-ABI helpers.
//
// Everything here is integer / file work that the reference also does on the host; the DP itself lives in sa_gpu.cu.
// Reference behaviour being re-implemented (paths relative to the signalAlign tree) is cited per function.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>
#include "sa_internal.h"

// ------------------------------------------------------------------------------------------------ errors
static thread_local char g_err[1024] = "";
void sa_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
extern "C" const char *sa_last_error(void) { return g_err; }

// ------------------------------------------------------------------------------------------------ params / ambiguity
extern "C" void sa_params_default(sa_params_t *p) {  // impl/pairwiseAligner.c:2022-2037
    p->threshold = 0.01;
    p->minDiagsBetweenTraceBack = 1000;
    p->traceBackDiagonals = 40;
    p->diagonalExpansion = 20;
    p->splitMatrixBiggerThanThis = (int64_t) 3000 * 3000;
}

void sa_ambig_default_table(sa_ambig &a) {  // create_ambig_bases, impl/pairwiseAligner.c:32-65
    memset(&a, 0, sizeof(a));
    static const char *tab[][2] = {
        {"R", "AG"}, {"Y", "CT"}, {"S", "CG"}, {"W", "AT"}, {"K", "GT"}, {"M", "AC"}, {"B", "CGT"}, {"D", "AGT"},
        {"H", "ACT"}, {"V", "ACG"}, {"X", "ACGT"}, {"L", "CEO"}, {"P", "CE"}, {"Q", "AI"}, {"f", "AF"},
        {"U", "ACEGOT"}, {"Z", "JT"}, {"j", "Tp"}, {"k", "Gb"}, {"l", "Gd"}, {"m", "Ce"}, {"n", "Th"}, {"o", "Ai"}};
    for (auto &t : tab) snprintf(a.opts[(unsigned char) t[0][0]], 16, "%s", t[1]);
}

extern "C" sa_ambig_t *sa_ambig_load(const char *file) {  // create_ambig_bases2, impl/pairwiseAligner.c:68-92
    sa_ambig *a = new sa_ambig;
    if (file == nullptr || file[0] == 0) {
        sa_ambig_default_table(*a);
        return a;
    }
    memset(a, 0, sizeof(*a));
    FILE *f = fopen(file, "r");
    if (!f) {
        sa_set_error("sa_ambig_load: cannot open %s", file);
        delete a;
        return nullptr;
    }
    char line[100], enc[100], bases[100];
    int i = 0;
    while (i < 300 && fgets(line, sizeof(line), f) != nullptr) {
        if (sscanf(line, "%s\t%s", enc, bases) == 2 && strlen(enc) == 1) {  // lookups are by single characters
            memset(a->opts[(unsigned char) enc[0]], 0, 16);
            snprintf(a->opts[(unsigned char) enc[0]], 16, "%s", bases);
        }
        i++;
    }
    fclose(f);
    return a;
}
extern "C" void sa_ambig_free(sa_ambig_t *a) { delete a; }

// An EMPTY table, filled entry by entry: what a caller holding the reference's stHash of code -> bases
// (Sequence.ambigBases, inc/pairwiseAligner.h:68) hands over; the drop-in shim converts the hash of every call this way.
extern "C" sa_ambig_t *sa_ambig_create(void) {
    sa_ambig *a = new sa_ambig;
    memset(a, 0, sizeof(*a));
    return a;
}
extern "C" int sa_ambig_set(sa_ambig_t *a, char code, const char *bases) {
    if (!a || !bases || strlen(bases) > 15) {
        sa_set_error("sa_ambig_set: a code stands for at most 15 bases");
        return SA_ERR_INPUT;
    }
    memset(a->opts[(unsigned char) code], 0, 16);
    snprintf(a->opts[(unsigned char) code], 16, "%s", bases);
    return SA_OK;
}

// ------------------------------------------------------------------------------------------------ kmers
static int64_t kmerId(const char *kmer, const sa_model &m) {  // kmer_id, impl/nanopore_hdp.c:377-410
    int64_t id = 0;
    for (int i = 0; i < m.k; i++) {
        size_t j = m.alphabet.find(kmer[i]);
        if (j == std::string::npos) return -1;
        id = id * m.A + (int64_t) j;
    }
    return id;
}

bool sa_kmer_paths(const char *kmer, const sa_model &m, const sa_ambig &a, std::vector<int32_t> &out) {
    // hdCell_construct2 (impl/pairwiseAligner.c:723-801): positions left to right, options in table order;
    // every existing kmer is expanded by every option, so the left-most ambiguous position is the most
    // significant digit of the path index.
    out.clear();
    int k = m.k;
    bool anyAmbig = false;
    for (int i = 0; i < k; i++)
        if (a.opts[(unsigned char) kmer[i]][0]) anyAmbig = true;
    if (!anyAmbig) {
        int64_t id = kmerId(kmer, m);
        if (id < 0) return false;
        out.push_back((int32_t) id);
        return true;
    }
    std::vector<std::string> list(1, std::string(kmer, k));
    for (int i = 0; i < k; i++) {
        const char *opts = a.opts[(unsigned char) kmer[i]];
        size_t no = strlen(opts);
        if (no == 0) continue;
        std::vector<std::string> nl;
        nl.reserve(list.size() * no);
        for (auto &s : list)
            for (size_t o = 0; o < no; o++) {
                std::string t = s;
                t[i] = opts[o];
                nl.push_back(t);
            }
        list.swap(nl);
    }
    for (auto &s : list) {
        int64_t id = kmerId(s.c_str(), m);
        if (id < 0) return false;
        out.push_back((int32_t) id);
    }
    return true;
}

extern "C" void sa_kmer_string(const sa_model_t *m, int64_t idx, char *out) {
    for (int i = m->k - 1; i >= 0; i--) {
        out[i] = idx >= 0 ? m->alphabet[idx % m->A] : '?';
        if (idx >= 0) idx /= m->A;
    }
    out[m->k] = 0;
}

// ------------------------------------------------------------------------------------------------ band
static inline int64_t dX(int64_t xay, int64_t xmy) { return (xay + xmy) / 2; }  // C truncating division, as reference
static inline int64_t dY(int64_t xay, int64_t xmy) { return (xay - xmy) / 2; }
static inline int64_t offByOne(int64_t xay, int64_t xmy) { return (xay + xmy) % 2 == 0 ? xmy : xmy + 1; }
static inline int64_t clampCoord(int64_t z, int64_t lZ) { return z < 0 ? 0 : (z > lZ ? lZ : z); }

void sa_build_band(const int64_t *anchors, int64_t nA, int64_t lX, int64_t lY, int64_t e, std::vector<Diag> &out) {
    // band_construct + band_setCurrentDiagonal, impl/pairwiseAligner.c:161-246
    out.resize(lX + lY + 1);
    int64_t ai = 0, xay = 0, pxay = 0, pxmy = 0, nxay = 0, nxmy = 0, xL = 0, yL = 0, xU = 0, yU = 0;
    while (xay <= lX + lY) {
        int64_t l = offByOne(xay, xL - yL), r = offByOne(xay, xU - yU);
        int64_t t;
        if ((t = xL - dX(xay, l)) > 0) l += 2 * t;
        if ((t = dY(xay, l) - yL) > 0) l += 2 * t;
        if ((t = dX(xay, r) - xU) > 0) r -= 2 * t;
        if ((t = yU - dY(xay, r)) > 0) r -= 2 * t;
        out[xay] = Diag{xay, l, r};
        if (nxay == xay++) {
            pxay = nxay;
            pxmy = nxmy;
            int64_t x = lX, y = lY;
            if (ai < nA) {
                x = anchors[2 * ai] + 1;
                y = anchors[2 * ai + 1] + 1;
                ai++;
            }
            nxay = x + y;
            nxmy = x - y;
            xL = clampCoord(dX(pxay, pxmy - e), lX);
            yL = clampCoord(dY(nxay, nxmy - e), lY);
            xU = clampCoord(dX(nxay, nxmy + e), lX);
            yU = clampCoord(dY(pxay, pxmy + e), lY);
        }
    }
}

// ------------------------------------------------------------------------------------------------ split
static bool splitP(int64_t &x1, int64_t &y1, int64_t x2, int64_t y2, int64_t x3, int64_t y3, std::vector<int64_t> &sp,
                   int64_t bigger, bool skip) {  // getSplitPointsP, impl/pairwiseAligner.c:1886-1908
    int64_t lX2 = x3 - x2, lY2 = y3 - y2;
    if (lX2 * lY2 > bigger) {
        int64_t maxLen = (int64_t) sqrt((double) bigger);
        int64_t hX = lX2 / 2 > maxLen ? maxLen : lX2 / 2;
        int64_t hY = lY2 / 2 > maxLen ? maxLen : lY2 / 2;
        if (!skip) {
            sp.push_back(x1); sp.push_back(y1); sp.push_back(x2 + hX); sp.push_back(y2 + hY);
        }
        x1 = x3 - hX;
        y1 = y3 - hY;
        return true;
    }
    return false;
}

void sa_split_points(const int64_t *anchors, int64_t nA, int64_t lX, int64_t lY, int64_t bigger, bool rl, bool rr,
                     std::vector<int64_t> &sp) {  // getSplitPoints, impl/pairwiseAligner.c:1910-1937
    sp.clear();
    int64_t x1 = 0, y1 = 0, x2 = 0, y2 = 0;
    for (int64_t i = 0; i < nA; i++) {
        int64_t x3 = anchors[2 * i], y3 = anchors[2 * i + 1];
        splitP(x1, y1, x2, y2, x3, y3, sp, bigger, rl && i == 0);
        x2 = x3 + 1;
        y2 = y3 + 1;
    }
    if (!splitP(x1, y1, x2, y2, lX, lY, sp, bigger, rl && nA == 0) || !rr) {
        sp.push_back(x1); sp.push_back(y1); sp.push_back(lX); sp.push_back(lY);
    }
}

// ------------------------------------------------------------------------------------------------ models
void sa_model_drop_device(const sa_model *m);  // sa_gpu.cu

static std::string sortedAlphabet(const char *alphabet) {  // sequence_prepareAlphabet, impl/pairwiseAligner.c:366-395
    std::string s(alphabet);
    std::sort(s.begin(), s.end());
    return s;
}

static bool readLine(FILE *f, std::string &out) {
    out.clear();
    int c;
    bool any = false;
    while ((c = fgetc(f)) != EOF) {
        any = true;
        if (c == '\n') break;
        out.push_back((char) c);
    }
    return any;
}

static void splitDoubles(const std::string &line, std::vector<double> &out) {
    out.clear();
    const char *p = line.c_str();
    char *end;
    for (;;) {
        while (*p == ' ' || *p == '\t' || *p == '\r') p++;
        if (!*p) break;
        double v = strtod(p, &end);
        if (end == p) break;
        out.push_back(v);
        p = end;
    }
}

extern "C" sa_model_t *sa_model_from_tables(const char *alphabet, int32_t k, const double *logTransitions,
                                            const double *matchTable, const double *gapYTable,
                                            const double *gapXTable, int kind) {
    sa_model *m = new sa_model;
    m->kind = kind;
    m->alphabet = sortedAlphabet(alphabet);
    m->A = (int) m->alphabet.size();
    m->k = k;
    m->K = 1;
    for (int i = 0; i < k; i++) m->K *= m->A;
    memcpy(m->trans, logTransitions, sizeof(double) * 9);
    m->match.assign(matchTable, matchTable + m->K * 5);
    if (gapYTable) {
        m->gapY.assign(gapYTable, gapYTable + m->K * 5);
    } else {  // EXTRA_EVENT_NOISE_MULTIPLIER, impl/stateMachine.c:1528-1532
        m->gapY = m->match;
        for (int64_t i = 0; i < m->K; i++) m->gapY[i * 5 + 1] *= 1.75;
    }
    if (gapXTable) m->gapX.assign(gapXTable, gapXTable + m->K);
    else m->gapX.assign(m->K, -2.3025850929940455);  // log(0.1), impl/stateMachine.c:1585
    return m;
}

static bool loadNhdp(sa_model *m, const char *path);

extern "C" sa_model_t *sa_model_load(const char *modelFile, const char *nhdpFile, int kind) {
    // stateMachine3_loadFromFile, impl/stateMachine.c:1440-1538
    FILE *f = fopen(modelFile, "r");
    if (!f) {
        sa_set_error("sa_model_load: cannot open model file %s", modelFile);
        return nullptr;
    }
    std::string l0, l1, l2;
    readLine(f, l0); readLine(f, l1); readLine(f, l2);
    fclose(f);
    long long stateNumber = 0, aSize = 0, kLen = 0;
    char alpha[256];
    if (sscanf(l0.c_str(), "%lld %lld %255s %lld", &stateNumber, &aSize, alpha, &kLen) != 4) {
        sa_set_error("sa_model_load: %s: malformed header line", modelFile);
        return nullptr;
    }
    if (stateNumber != 3 || (long long) strlen(alpha) != aSize) {
        sa_set_error("sa_model_load: %s: need a 3-state model with alphabetSize == strlen(alphabet)", modelFile);
        return nullptr;
    }
    std::vector<double> t, e;
    splitDoubles(l1, t);
    if (t.size() != 10) {
        sa_set_error("sa_model_load: %s: transitions line has %zu tokens, expected 10", modelFile, t.size());
        return nullptr;
    }
    splitDoubles(l2, e);
    int64_t K = 1;
    for (int i = 0; i < kLen; i++) K *= aSize;
    if ((int64_t) e.size() != K * 5) {
        sa_set_error("sa_model_load: %s: emissions line has %zu tokens, expected %lld", modelFile, e.size(),
                     (long long) K * 5);
        return nullptr;
    }
    // stateMachine3_loadTransitionsFromFile (impl/stateMachine.c:1202-1258): file order is row-major M,X,Y x M,X,Y;
    // token 5 (gapX->gapY) is skipped, token 7 (gapY->gapX) lands in SWITCH_TO_Y, SWITCH_TO_X keeps its -inf default.
    double lt[9];
    lt[T_MATCH_CONTINUE] = log(t[0]);
    lt[T_GAP_OPEN_X] = log(t[1]);
    lt[T_GAP_OPEN_Y] = log(t[2]);
    lt[T_MATCH_FROM_GAP_X] = log(t[3]);
    lt[T_GAP_EXTEND_X] = log(t[4]);
    lt[T_MATCH_FROM_GAP_Y] = log(t[6]);
    lt[T_GAP_SWITCH_TO_Y] = log(t[7]);
    lt[T_GAP_EXTEND_Y] = log(t[8]);
    lt[T_GAP_SWITCH_TO_X] = -INFINITY;
    sa_model *m = sa_model_from_tables(alpha, (int32_t) kLen, lt, e.data(), nullptr, nullptr, kind);
    if (nhdpFile && nhdpFile[0]) {
        if (!loadNhdp(m, nhdpFile)) {
            delete m;
            return nullptr;
        }
        m->kind = SA_EMIT_HDP;
    }
    return m;
}

static bool loadNhdp(sa_model *m, const char *path) {
    // deserialize_nhdp (impl/nanopore_hdp.c:1088-1115) + deserialize_hdp (impl/hdp.c:3052-3322): only what the
    // density query needs (dir_proc_density, impl/hdp.c:2588-2612): grid, parents, observed flags, posterior
    // predictive and spline slopes of observed DPs.
    FILE *f = fopen(path, "r");
    if (!f) {
        sa_set_error("sa_model_load: cannot open nhdp file %s", path);
        return false;
    }
    std::string line;
    auto fail = [&](const char *what) {
        sa_set_error("sa_model_load: %s: %s", path, what);
        fclose(f);
        return false;
    };
    long long aSize, kLen;
    char alpha[256];
    readLine(f, line); aSize = atoll(line.c_str());
    readLine(f, line); if (sscanf(line.c_str(), "%255s", alpha) != 1) return fail("bad alphabet");
    readLine(f, line); kLen = atoll(line.c_str());
    if (sortedAlphabet(alpha) != m->alphabet || kLen != m->k || aSize != m->A)
        return fail("alphabet / kmer length differ from the .model file");
    readLine(f, line); bool splines = atol(line.c_str()) != 0;
    readLine(f, line); bool hasData = atol(line.c_str()) != 0;
    readLine(f, line); bool sampleGamma = atol(line.c_str()) != 0;
    readLine(f, line); long long numDps = atoll(line.c_str());
    if (!splines) return fail("splines not finalized (the reference refuses to query such an HDP)");
    if (!hasData) return fail("HDP without data has no observed distributions");
    std::vector<double> tmp;
    std::vector<long long> dpIds;
    readLine(f, line);  // data values (unused)
    readLine(f, line);  // data point dp ids
    {
        const char *p = line.c_str();
        char *end;
        for (;;) {
            while (*p == ' ' || *p == '\t') p++;
            if (!*p) break;
            long long v = strtoll(p, &end, 10);
            if (end == p) break;
            dpIds.push_back(v);
            p = end;
        }
    }
    readLine(f, line);  // mu nu alpha beta
    readLine(f, line);
    double gStart, gStop;
    long long G;
    if (sscanf(line.c_str(), "%lg\t%lg\t%lld", &gStart, &gStop, &G) != 3) return fail("bad grid line");
    readLine(f, line);  // gamma
    if (sampleGamma) { readLine(f, line); readLine(f, line); readLine(f, line); readLine(f, line); }
    std::vector<long long> parent(numDps, -1);
    for (long long id = 0; id < numDps; id++) {
        readLine(f, line);
        if (line.empty()) return fail("truncated parent table");
        if (line[0] != '-') parent[id] = atoll(line.c_str());
    }
    if (numDps < m->K) return fail("fewer Dirichlet processes than kmers");
    // mark_observed_dps (impl/hdp.c:1132-1160)
    std::vector<char> observed(numDps, 0);
    for (long long id : dpIds) {
        if (id < 0 || id >= numDps) return fail("data point assigned to a missing DP");
        long long d = id;
        while (d >= 0 && !observed[d]) {
            observed[d] = 1;
            d = parent[d];
        }
    }
    // linspace (impl/hdp_math_utils.c:497-510)
    m->G = G;
    m->grid.resize(G);
    {
        double dx = (gStop - gStart) / ((double) (G - 1));
        for (long long i = 0; i < G - 1; i++) m->grid[i] = gStart + i * dx;
        m->grid[G - 1] = gStop;
    }
    std::vector<int32_t> rowOf(numDps, -1);
    m->nRows = 0;
    for (long long id = 0; id < numDps; id++)
        if (observed[id]) rowOf[id] = (int32_t) m->nRows++;
    m->postPred.assign(m->nRows * G, 0.0);
    m->slopes.assign(m->nRows * G, 0.0);
    for (int pass = 0; pass < 2; pass++) {
        std::vector<double> &dst = pass == 0 ? m->postPred : m->slopes;
        for (long long id = 0; id < numDps; id++) {
            if (!readLine(f, line)) return fail("truncated distribution table");
            splitDoubles(line, tmp);
            if (tmp.empty()) continue;
            if ((long long) tmp.size() < G) return fail("short distribution line");
            if (rowOf[id] < 0) continue;  // the reference would store it, but never reads it (not observed)
            std::copy(tmp.begin(), tmp.begin() + G, dst.begin() + (int64_t) rowOf[id] * G);
        }
    }
    fclose(f);
    m->kmerToRow.assign(m->K, -1);
    m->ownRow.assign(m->K, -1);
    for (int64_t kI = 0; kI < m->K; kI++) {
        long long d = kI;
        m->ownRow[kI] = rowOf[kI];
        while (d >= 0 && !observed[d]) d = parent[d];  // dir_proc_density's parent walk
        if (d < 0) {
            sa_set_error("sa_model_load: %s: kmer %lld has no observed ancestor", path, (long long) kI);
            return false;
        }
        m->kmerToRow[kI] = rowOf[d];
    }
    return true;
}

extern "C" int sa_model_attach_nhdp(sa_model_t *m, const char *nhdpFile) {
    if (!m || !nhdpFile || !nhdpFile[0]) {
        sa_set_error("sa_model_attach_nhdp: no model or no file");
        return SA_ERR_INPUT;
    }
    if (!loadNhdp(m, nhdpFile)) return SA_ERR_INPUT;
    m->kind = SA_EMIT_HDP;
    sa_model_drop_device(m);
    return SA_OK;
}

extern "C" int sa_model_set_hdp(sa_model_t *m, int64_t G, const double *grid, const int32_t *kmerToRow,
                                const int32_t *ownRow, int64_t nRows, const double *postPred, const double *slopes) {
    m->kind = SA_EMIT_HDP;
    m->G = G;
    m->grid.assign(grid, grid + G);
    m->kmerToRow.assign(kmerToRow, kmerToRow + m->K);
    if (ownRow) m->ownRow.assign(ownRow, ownRow + m->K);
    else m->ownRow = m->kmerToRow;
    m->nRows = nRows;
    m->postPred.assign(postPred, postPred + nRows * G);
    m->slopes.assign(slopes, slopes + nRows * G);
    sa_model_drop_device(m);
    return SA_OK;
}

extern "C" int sa_model_set_hdp_expected_values(sa_model_t *m) {
    // stateMachine3_setModelToHdpExpectedValues (impl/stateMachine.c:1275-1304) with dir_proc_expected_val /
    // dir_proc_variance (impl/hdp.c:2777-2807): only kmers whose own DP is observed are replaced.
    if (m->G == 0) {
        sa_set_error("sa_model_set_hdp_expected_values: model has no HDP");
        return SA_ERR_INPUT;
    }
    for (int64_t kI = 0; kI < m->K; kI++) {
        int32_t row = m->ownRow[kI];
        if (row < 0) continue;
        const double *distr = m->postPred.data() + (int64_t) row * m->G;
        double ev = 0.0;
        for (int64_t i = 1; i < m->G; i++) {
            double dx = m->grid[i] - m->grid[i - 1];
            ev += m->grid[i] * distr[i] * dx;
        }
        double var = 0.0;
        for (int64_t i = 1; i < m->G; i++) {
            double dx = m->grid[i] - m->grid[i - 1];
            double dev = m->grid[i] - ev;
            var += dev * dev * distr[i] * dx;
        }
        m->match[kI * 5] = ev;
        m->match[kI * 5 + 1] = sqrt(var);
    }
    sa_model_drop_device(m);
    return SA_OK;
}

extern "C" int sa_model_scale_noise(sa_model_t *m, double scale_sd, double var_sd) {
    // emissions_signal_scaleNoise, impl/stateMachine.c:721-741 (note: the gap-Y noise sd divides by the MATCH lambda)
    for (int64_t i = 0; i < m->K * 5; i += 5) {
        m->match[i + 2] = m->match[i + 2] * scale_sd;
        m->match[i + 4] = m->match[i + 4] * var_sd;
        m->match[i + 3] = sqrt(pow(m->match[i + 2], 3.0) / m->match[i + 4]);
        m->gapY[i + 2] = m->gapY[i + 2] * scale_sd;
        m->gapY[i + 4] = m->gapY[i + 4] * var_sd;
        m->gapY[i + 3] = sqrt(pow(m->gapY[i + 2], 3.0) / m->match[i + 4]);
    }
    sa_model_drop_device(m);
    return SA_OK;
}

extern "C" int sa_model_scale_model(sa_model_t *m, double scale, double shift, double var, double scale_sd,
                                    double var_sd) {  // emissions_signal_scaleModel, impl/stateMachine.c:743-779
    for (int64_t i = 0; i < m->K * 5; i += 5) {
        m->match[i] = m->match[i] * scale + shift;
        m->match[i + 1] = m->match[i + 1] * var;
        m->gapY[i] = m->gapY[i] * scale + shift;
        m->gapY[i + 1] = m->gapY[i + 1] * var;
    }
    return sa_model_scale_noise(m, scale_sd, var_sd);
}

extern "C" void sa_model_free(sa_model_t *m) {
    if (!m) return;
    sa_model_drop_device(m);
    delete m;
}
extern "C" int64_t sa_model_num_kmers(const sa_model_t *m) { return m->K; }
extern "C" int32_t sa_model_kmer_length(const sa_model_t *m) { return m->k; }
extern "C" const char *sa_model_alphabet(const sa_model_t *m) { return m->alphabet.c_str(); }
extern "C" void sa_model_get_tables(const sa_model_t *m, double *t9, double *match, double *gapY) {
    if (t9) memcpy(t9, m->trans, sizeof(double) * 9);
    if (match) memcpy(match, m->match.data(), sizeof(double) * m->K * 5);
    if (gapY) memcpy(gapY, m->gapY.data(), sizeof(double) * m->K * 5);
}

extern "C" double sa_model_hdp_density(const sa_model_t *m, int64_t kI, double q) {
    // dir_proc_density (impl/hdp.c:2588-2612) + grid_spline_interp (impl/hdp_math_utils.c:471-495)
    if (m->G == 0 || kI < 0 || kI >= m->K) return NAN;
    int64_t n = m->G;
    const double *x = m->grid.data();
    const double *y = m->postPred.data() + (int64_t) m->kmerToRow[kI] * n;
    const double *s = m->slopes.data() + (int64_t) m->kmerToRow[kI] * n;
    double v;
    if (q <= x[0]) v = y[0] - s[0] * (x[0] - q);
    else if (q >= x[n - 1]) v = y[n - 1] + s[n - 1] * (q - x[n - 1]);
    else {
        double dx = x[1] - x[0];
        int64_t il = (int64_t) ((q - x[0]) / dx), ir = il + 1;
        double dy = y[ir] - y[il];
        double a = s[il] * dx - dy, b = dy - s[ir] * dx;
        double tl = (q - x[il]) / dx, tr = 1.0 - tl;
        v = tr * y[il] + tl * y[ir] + tl * tr * (a * tr + b * tl);
    }
    return v > 0.0 ? v : 0.0;
}

bool sa_pinned_release(void *p);
extern "C" void sa_result_free(sa_result_t *r) {
    if (!r) return;
    if (!sa_pinned_release(r->pairs)) free(r->pairs);   // normally a slab of the pinned pool (sa_gpu.cu)
    free(r->pairOffsets); free(r->transitions); free(r->likelihood);
    free(r->assignOffsets); free(r->assignKmerPos); free(r->assignEvent);
    free(r);
}

extern "C" void sa_result_sum_expectations(const sa_result_t *r, double *out10) {
    for (int i = 0; i < 10; i++) out10[i] = 0.0;
    if (!r->transitions) return;
    for (int64_t i = 0; i < r->nReads; i++) {
        for (int j = 0; j < 9; j++) out10[j] += r->transitions[i * 9 + j];
        out10[9] += r->likelihood[i];
    }
}

extern "C" void sa_model_hdp_dims(const sa_model_t *m, int64_t *dims2) {
    dims2[0] = m->G;
    dims2[1] = m->nRows;
}
extern "C" void sa_model_get_hdp(const sa_model_t *m, double *grid, int32_t *kmerToRow, int32_t *ownRow, double *postPred,
                                 double *slopes) {
    if (ownRow) memcpy(ownRow, m->ownRow.data(), sizeof(int32_t) * m->ownRow.size());
    if (grid) memcpy(grid, m->grid.data(), sizeof(double) * m->grid.size());
    if (kmerToRow) memcpy(kmerToRow, m->kmerToRow.data(), sizeof(int32_t) * m->kmerToRow.size());
    if (postPred) memcpy(postPred, m->postPred.data(), sizeof(double) * m->postPred.size());
    if (slopes) memcpy(slopes, m->slopes.data(), sizeof(double) * m->slopes.size());
}

// ------------------------------------------------------------------------------------------------ test hooks
extern "C" void sa_debug_band(const int64_t *anchors, int64_t nA, int64_t lX, int64_t lY, int64_t e, int64_t *out) {
    std::vector<Diag> band;
    sa_build_band(anchors, nA, lX, lY, e, band);
    for (size_t i = 0; i < band.size(); i++) {
        out[3 * i] = band[i].xay; out[3 * i + 1] = band[i].xmyL; out[3 * i + 2] = band[i].xmyR;
    }
}
extern "C" int64_t sa_debug_split_points(const int64_t *anchors, int64_t nA, int64_t lX, int64_t lY, int64_t bigger,
                                         int rl, int rr, int64_t *out, int64_t cap) {
    std::vector<int64_t> sp;
    sa_split_points(anchors, nA, lX, lY, bigger, rl != 0, rr != 0, sp);
    int64_t n = (int64_t) sp.size() / 4;
    for (int64_t i = 0; i < n && i < cap; i++)
        for (int j = 0; j < 4; j++) out[4 * i + j] = sp[4 * i + j];
    return n;
}
extern "C" int64_t sa_debug_kmer_paths(const sa_model_t *m, const sa_ambig_t *ambig, const char *kmer, int32_t *out,
                                       int64_t cap) {
    sa_ambig def;
    if (!ambig) {
        sa_ambig_default_table(def);
        ambig = &def;
    }
    std::vector<int32_t> paths;
    if (!sa_kmer_paths(kmer, *m, *ambig, paths)) return -1;
    for (int64_t i = 0; i < (int64_t) paths.size() && i < cap; i++) out[i] = paths[i];
    return (int64_t) paths.size();
}
