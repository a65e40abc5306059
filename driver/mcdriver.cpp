// mcdriver — C++ stand-in for the reference's Fortran driver on the classical LJ path (phase 1 only).
//
// The reference's driver (MonteCarlo/MonteCarlo.F90) stays Fortran and keeps its config and file formats; no
// Fortran compiler exists in the build image, so this program plays its part above the C ABI: same INI keys
// (readConfig.f90:89-161), same box / DMSO / param / par_solute readers (MonteCarlo.F90:1053-1166), same OUT rows
// (format 902, :393/:762), DUMP frames (:1009-1019), `solventsolvent` matrix (:372-376), result box (:587-611)
// and stdout energy lines (:367, :437).  MCINIT (:685-715) and METROPOLIS (:720-788) run here on the host, as
// they would in the Fortran driver; the energies come from libmcb200.so (mcb_total_energy, mcb_trial_delta,
// mcb_commit).  With `mode=batched` the whole loop_LJ runs on the GPU through mcb_run_steps instead.
//
// Differences from the reference, on purpose: `seed` is honoured (the reference ignores it, randgen.f90:36-40) and
// feeds the library's documented counter stream; GA (QM) steps are not available; Gaussian is never called.
//
//   mcdriver <config.ini> [mode=single|batched] [chains=N]
#include "../include/mcb200.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

namespace {

struct Config {
    long lj = 0, ga = 0, prod = 1, seed = 1, lj_nadj = 100, lj_nprint = 0, lj_dump = 0, proc = 1;
    double temp = 300.0, dposmax = 0.3, dposmin = 1e-10, dhoekmax = 0.16, dhoekmin = 1e-10, padj = 0.05, drotsolu = 60.0;
    std::string box, dmso, solute, param, out, err, dump, solvsolv, result, parsol, solout;
};

[[noreturn]] void die(const std::string &m) { fprintf(stderr, "mcdriver: %s\n", m.c_str()); exit(2); }

std::string lower(std::string s) { for (auto &c : s) c = (char)tolower(c); return s; }

double fnum(std::string t) { for (auto &c : t) if (c == 'D' || c == 'd') c = 'E'; return strtod(t.c_str(), nullptr); }

// rConfig, readConfig.f90:53-163: `key = value`, '#'/';' comments, case-insensitive keys, unknown key -> stop
Config read_config(const std::string &path)
{
    std::ifstream f(path);
    if (!f) die("cannot open config " + path);
    Config c;
    std::string line;
    while (std::getline(f, line)) {
        if (line.empty() || line[0] == '#' || line[0] == ';') continue;
        size_t b = line.find_first_not_of(" \t");
        if (b == std::string::npos) continue;
        size_t e = line.find_first_of("= \t", b);
        std::string key = lower(line.substr(b, e == std::string::npos ? std::string::npos : e - b)), val;
        if (e != std::string::npos) { size_t v = line.find_first_not_of("= \t", e); if (v != std::string::npos) val = line.substr(v); }
        while (!val.empty() && (val.back() == ' ' || val.back() == '\r' || val.back() == '\t')) val.pop_back();
        if (key == "lj") c.lj = atol(val.c_str()); else if (key == "ga") c.ga = atol(val.c_str());
        else if (key == "prod") c.prod = atol(val.c_str()); else if (key == "temp") c.temp = fnum(val);
        else if (key == "seed") c.seed = atol(val.c_str()); else if (key == "lj_nadj") c.lj_nadj = atol(val.c_str());
        else if (key == "lj_nprint") c.lj_nprint = atol(val.c_str()); else if (key == "lj_dump") c.lj_dump = atol(val.c_str());
        else if (key == "ga_nadj" || key == "ga_nprint" || key == "ga_dump" || key == "rotsolu") {}
        else if (key == "dposmax") c.dposmax = fnum(val); else if (key == "dposmin") c.dposmin = fnum(val);
        else if (key == "dhoekmax") c.dhoekmax = fnum(val); else if (key == "dhoekmin") c.dhoekmin = fnum(val);
        else if (key == "padj") c.padj = fnum(val); else if (key == "proc") c.proc = atol(val.c_str());
        else if (key == "drotsolu") c.drotsolu = fnum(val);
        else if (key == "box") c.box = val; else if (key == "dmso") c.dmso = val; else if (key == "solute") c.solute = val;
        else if (key == "param") c.param = val; else if (key == "out") c.out = val; else if (key == "err") c.err = val;
        else if (key == "dump") c.dump = val; else if (key == "solventsolvent") c.solvsolv = val;
        else if (key == "result") c.result = val; else if (key == "parsol") c.parsol = val; else if (key == "solout") c.solout = val;
        else die("unknown option '" + key + "'");   // readConfig.f90:159-160
    }
    return c;
}

std::vector<std::string> read_lines(const std::string &p)
{
    std::ifstream f(p);
    if (!f) die("cannot open " + p);
    std::vector<std::string> v;
    std::string l;
    while (std::getline(f, l)) v.push_back(l);
    return v;
}

std::vector<std::string> toks(const std::string &l) { std::istringstream s(l); std::vector<std::string> t; std::string x; while (s >> x) t.push_back(x); return t; }

struct Box {
    double boxl = 0, energy = 123456.789, e_sol = 0;
    std::string id, sol_name;
    std::vector<double> com, hoek, solute;     // [n][3]
    std::vector<std::string> sol_sym;
    std::vector<std::string> dih_lines;
    int n = 0, nsol = 0;
};

// read_norm, MonteCarlo.F90:1078-1128
Box read_box(const std::string &path)
{
    auto L = read_lines(path);
    Box b;
    size_t p = 0;
    b.boxl = fnum(toks(L.at(p++)).at(0));
    b.n = atoi(toks(L.at(p++)).at(0).c_str());
    { auto t = toks(L.at(p)); char *end = nullptr; std::string s = t.empty() ? "" : t[0]; for (auto &c : s) if (c == 'D') c = 'E';
      double v = strtod(s.c_str(), &end); if (!s.empty() && end && *end == 0) { b.energy = v; ++p; } }
    b.id = L.at(p++);
    for (int i = 0; i < b.n; ++i) { auto t = toks(L.at(p++)); for (int k = 0; k < 3; ++k) { b.com.push_back(fnum(t.at(k))); b.hoek.push_back(fnum(t.at(3 + k))); } }
    ++p;   // SOLUTE
    b.nsol = atoi(toks(L.at(p++)).at(0).c_str());
    b.sol_name = L.at(p++);
    for (int i = 0; i < b.nsol; ++i) { auto t = toks(L.at(p++)); b.sol_sym.push_back(t.at(0)); for (int k = 0; k < 3; ++k) b.solute.push_back(fnum(t.at(1 + k))); }
    b.e_sol = fnum(toks(L.at(p++)).at(0));
    int nd = atoi(toks(L.at(p++)).at(0).c_str());
    for (int i = 0; i < nd; ++i) b.dih_lines.push_back(L.at(p++));
    return b;
}

int find_sym(const std::string &s, const std::vector<std::string> &tab) { for (size_t i = 0; i < tab.size(); ++i) if (tab[i] == s) return (int)i; return 0; }   // findSym, lib.f90:87-107

// F6.5 of the reference's format 902 has no room for a leading zero: ".30000"
std::string f65(double v) { char b[32]; snprintf(b, sizeof b, "%.5f", v); std::string s = b; if (s.rfind("0.", 0) == 0) s = s.substr(1); if (s.size() > 6) s = "******"; while (s.size() < 6) s = " " + s; return s; }

inline uint64_t mix64(uint64_t x) { x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL; x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL; return x ^ (x >> 31); }
inline double stream_u(uint64_t key, uint64_t n) { return (double)(mix64(key + (n + 1) * 0x9E3779B97F4A7C15ULL) >> 11) * 0x1.0p-53; }

void ck(mcb_t *h, int rc, const char *what) { if (rc) die(std::string(what) + " failed: " + mcb_last_error(h)); }

void dump_frame(FILE *f, const Box &b, long step)   // dump, MonteCarlo.F90:1009-1019
{
    fprintf(f, " %11d\n Timestep: %11ld\n", b.n, step);
    for (int j = 0; j < b.n; ++j)
        fprintf(f, " %.17g %.17g %.17g %.17g %.17g %.17g\n", b.com[3 * j], b.com[3 * j + 1], b.com[3 * j + 2], b.hoek[3 * j], b.hoek[3 * j + 1], b.hoek[3 * j + 2]);
}

}  // namespace

int main(int argc, char **argv)
{
    if (argc < 2) die("usage: mcdriver <config.ini> [mode=single|batched] [chains=N]");
    std::string mode = "single";
    int chains = 1;
    for (int a = 2; a < argc; ++a) {
        if (!strncmp(argv[a], "mode=", 5)) mode = argv[a] + 5;
        else if (!strncmp(argv[a], "chains=", 7)) chains = atoi(argv[a] + 7);
        else die(std::string("unknown argument ") + argv[a]);
    }
    Config cfg = read_config(argv[1]);
    const double PI = 4.0 * atan(1.0), TAU = 2.0 * PI;
    cfg.dhoekmax *= PI; cfg.dhoekmin *= PI;                               // MonteCarlo.F90:219-220
    long start_prod = std::max(0L, cfg.lj + cfg.ga - cfg.prod + 1);       // :222
    if (cfg.lj_dump == 0) cfg.lj_dump = 0x7fffffff;
    const bool do_out = cfg.lj_nprint != 0, do_dump = cfg.lj_dump != 0x7fffffff;
    if (cfg.lj_nprint == 0) cfg.lj_nprint = 0x7fffffff;

    // ---- inputs (read_norm)
    Box box = read_box(cfg.box);
    auto D = read_lines(cfg.dmso);
    int nd = atoi(toks(D.at(0)).at(0).c_str());
    std::vector<std::string> dsym; std::vector<double> drel;
    for (int i = 0; i < nd; ++i) { auto t = toks(D.at(2 + i)); dsym.push_back(t.at(0)); for (int k = 0; k < 3; ++k) drel.push_back(fnum(t.at(1 + k))); }
    auto P = read_lines(cfg.param);
    int np = atoi(toks(P.at(0)).at(0).c_str());
    std::vector<std::string> psym; std::vector<double> ptab;
    for (int i = 0; i < np; ++i) { auto t = toks(P.at(2 + i)); psym.push_back(t.at(0)); for (int k = 0; k < 3; ++k) ptab.push_back(fnum(t.at(1 + k))); }
    auto S = read_lines(cfg.parsol);
    int ns = atoi(toks(S.at(0)).at(0).c_str());
    std::vector<std::string> ssym; std::vector<double> stab;
    for (int i = 0; i < ns; ++i) { auto t = toks(S.at(2 + i)); ssym.push_back(t.at(0)); stab.push_back(fnum(t.at(1))); stab.push_back(fnum(t.at(2))); }
    // ASSOSIATE_DMSO / ASSOSIATE_SOLUTE (LennardJones.f90:192-248); only the non-H sites go over the ABI (:66,68)
    std::vector<double> site_rel, site_tab, sol_tab;
    for (int i = 0; i < nd; ++i) if (dsym[i] != "H") {
        int a = find_sym(dsym[i], psym);
        for (int k = 0; k < 3; ++k) { site_rel.push_back(drel[3 * i + k]); site_tab.push_back(ptab[3 * a + k]); }
    }
    for (int i = 0; i < box.nsol; ++i) { int a = find_sym(box.sol_sym[i], ssym); sol_tab.push_back(0.0); sol_tab.push_back(stab[2 * a]); sol_tab.push_back(stab[2 * a + 1]); }

    mcb_params prm{};
    prm.n_mol = box.n; prm.n_sites = (int)site_rel.size() / 3; prm.n_sol = box.nsol; prm.n_chains = chains; prm.device = -1;
    prm.quirk_31 = 1; prm.boxl = box.boxl; prm.cutoff = 7.0; prm.conv = 0.0;   // CONV == 0: LennardJones.f90:57
    prm.site_rel = site_rel.data(); prm.site_tab = site_tab.data();
    mcb_t *h = nullptr;
    if (int rc = mcb_create(&h, &prm)) die(std::string("mcb_create failed (") + std::to_string(rc) + "): " + mcb_last_error(nullptr));
    if (box.nsol) ck(h, mcb_set_solute(h, 0, chains, box.solute.data(), sol_tab.data(), 1), "mcb_set_solute");
    ck(h, mcb_set_chain_state(h, 0, chains, box.com.data(), box.hoek.data(), 1), "mcb_set_chain_state");

    // ---- start-up energy (:357-376)
    std::vector<double> tot(chains);
    ck(h, mcb_total_energy(h, 0, chains, tot.data(), nullptr, nullptr), "mcb_total_energy");
    double toteng = tot[0];
    printf("Initial energy (LJ): %20.10E\n", toteng);
    if (!cfg.solvsolv.empty()) {
        std::vector<double> ss((size_t)box.n * box.n);
        ck(h, mcb_get_matrix(h, 0, ss.data(), nullptr), "mcb_get_matrix");
        FILE *f = fopen(cfg.solvsolv.c_str(), "w");
        if (f) { for (int i = 0; i < box.n; ++i) { for (int j = 0; j < box.n; ++j) fprintf(f, " %.15g", ss[(size_t)i * box.n + j]); fprintf(f, "\n"); } fclose(f); }
    }
    FILE *fout = do_out && !cfg.out.empty() ? fopen(cfg.out.c_str(), "w") : nullptr;
    FILE *fdump = do_dump && !cfg.dump.empty() ? fopen(cfg.dump.c_str(), "w") : nullptr;
    if (fdump) { fprintf(fdump, " %.17g\n", box.boxl); dump_frame(fdump, box, 0); }
    if (fout) {
        fprintf(fout, "%12s %20s %20s %6s %6s %3s %6s %6s %6s %6s\n", "i", "TotEng", "TotEng_old", "kans", "rv", "rSo", "pSuc", "ratio", "dposma", "dhoekm");
        fprintf(fout, "%012d %20.10E %20.10E %6.4f %6.4f %03d %6.4f %6.4f %s %s\n", 0, toteng, toteng, 0.0, 0.0, 0, 0.0, 0.0, f65(cfg.dposmax).c_str(), f65(cfg.dhoekmax).c_str());
    }

    double toteng_old = toteng, production = 0;
    long nsuc = 0, prod_acc = 0;
    if (mode == "batched") {
        // ---- loop_LJ on the GPU for every chain (:406-433)
        mcb_mc_params mp{cfg.temp, cfg.dposmax, cfg.dposmin, cfg.dhoekmax, cfg.dhoekmin, cfg.padj, (int)cfg.lj_nadj, 0, start_prod};
        ck(h, mcb_set_mc_params(h, 0, chains, &mp, 1), "mcb_set_mc_params");
        std::vector<uint64_t> seeds(chains);
        for (int c = 0; c < chains; ++c) seeds[c] = (uint64_t)cfg.seed + c;
        ck(h, mcb_seed_chains(h, 0, chains, seeds.data()), "mcb_seed_chains");
        ck(h, mcb_run_steps(h, cfg.lj, nullptr, 0, nullptr, nullptr), "mcb_run_steps");
        std::vector<mcb_chain_stats> st(chains);
        ck(h, mcb_get_stats(h, 0, chains, st.data()), "mcb_get_stats");
        ck(h, mcb_get_chain_state(h, 0, 1, box.com.data(), box.hoek.data()), "mcb_get_chain_state");
        toteng_old = toteng = st[0].toteng; nsuc = st[0].nsuc; production = st[0].production; prod_acc = st[0].prod_acc;
        for (int c = 0; c < chains; ++c)
            printf("chain %d: steps %lld accepted %lld energy %20.10E dposmax %.5f site-pairs %lld\n", c, (long long)st[c].step,
                   (long long)st[c].nsuc, st[c].toteng, st[c].dposmax, (long long)st[c].sitepairs);
    } else {
        // ---- loop_LJ on the host, one chain, energies from the GPU (single-step mode)
        const uint64_t key = mix64((uint64_t)cfg.seed + 0x9E3779B97F4A7C15ULL);
        const double E = (double)2.71828182845904523536f;               // lib.f90:17
        const double inv_sqrt3 = 1.0 / sqrt(3.0);
        double dposmax = cfg.dposmax, dhoekmax = cfg.dhoekmax;
        long naccept = 0;
        const double boxl = box.boxl, boxl2 = 2.0 * box.boxl;
        for (long step = 1; step <= cfg.lj; ++step) {
            const uint64_t base = 8ULL * (uint64_t)(step - 1);
            // MCINIT (:685-715)
            toteng_old = toteng;
            int rsolv = (int)(stream_u(key, base) * box.n) + 1;
            if (rsolv == 31) rsolv = 30;
            const int i = rsolv - 1;
            double oc[3], oh[3], nc[3], nh[3];
            for (int k = 0; k < 3; ++k) { oc[k] = box.com[3 * i + k]; oh[k] = box.hoek[3 * i + k]; }
            for (int k = 0; k < 3; ++k) { double rv = (stream_u(key, base + 1 + k) - 0.5) * 2; rv = inv_sqrt3 * rv; rv = dposmax * rv; nc[k] = oc[k] + rv; }
            for (int k = 0; k < 3; ++k) nh[k] = oh[k] + (stream_u(key, base + 4 + k) - 0.5) * 2 * dhoekmax;
            for (int k = 0; k < 3; ++k) nh[k] = nh[k] - TAU * trunc(nh[k] / PI);
            for (int k = 0; k < 3; ++k) nc[k] = nc[k] - boxl2 * trunc(nc[k] / boxl);
            // calculateLJ(RSOLV) + calcEnergy (:411-414)
            double d_e = 0, e_new = 0;
            ck(h, mcb_trial_delta(h, 1, &rsolv, nc, nh, &d_e, &e_new), "mcb_trial_delta");
            toteng = e_new;
            for (int k = 0; k < 3; ++k) { box.com[3 * i + k] = nc[k]; box.hoek[3 * i + k] = nh[k]; }
            if (fdump && step % cfg.lj_dump == 0) dump_frame(fdump, box, step);          // the TRIAL configuration (:416-418)
            if (!(toteng <= 1.7976931348623157e308)) toteng = 1.7976931348623157e308;    // :421-428
            // METROPOLIS (:720-788)
            const double delta = toteng - toteng_old;
            const double exponent = -1.0 * delta * 1000.0 / (8.314 * cfg.temp);
            double kans, rv;
            if (exponent < -75.0) { kans = 0; rv = 1; }
            else if (exponent >= 0) { kans = 1; rv = 0; }
            else { kans = std::min(pow(E, exponent), 1.0); rv = stream_u(key, base + 7); }
            const bool accepted = rv <= kans;
            if (accepted) { ++nsuc; ++naccept; if (step >= start_prod) { production += toteng; ++prod_acc; } }
            const double ratio = (double)((float)naccept / (float)cfg.lj_nadj);
            if (fout && step % cfg.lj_nprint == 0)
                fprintf(fout, "%012ld %20.10E %20.10E %6.4f %6.4f %03d %6.4f %6.4f %s %s\n", step, toteng, toteng_old, kans, rv, rsolv,
                        (double)((float)nsuc / (float)step), ratio, f65(dposmax).c_str(), f65(dhoekmax).c_str());
            if (step % cfg.lj_nadj == 0) {
                if (ratio > 0.5) { dposmax *= 1.0 + cfg.padj; dhoekmax *= 1.0 + cfg.padj; }
                else if (ratio < 0.5) { dposmax *= 1.0 - cfg.padj; dhoekmax *= 1.0 - cfg.padj; }
                dposmax = std::max(dposmax, cfg.dposmin); dhoekmax = std::max(dhoekmax, cfg.dhoekmin);
                naccept = 0;
            }
            unsigned char acc = accepted ? 1 : 0;
            ck(h, mcb_commit(h, 1, &acc), "mcb_commit");
            if (!accepted) { for (int k = 0; k < 3; ++k) { box.com[3 * i + k] = oc[k]; box.hoek[3 * i + k] = oh[k]; } toteng = toteng_old; }
        }
    }
    if (cfg.lj > 0) printf("Final energy (LJ): %20.10E\n", toteng_old);                  // :437 prints TOTENG_OLD
    if (fdump) { dump_frame(fdump, box, cfg.lj + 2); fclose(fdump); }                     // :581
    if (fout) fclose(fout);
    printf("Production: %ld accepted steps, mean %20.10E\n", prod_acc, prod_acc ? production / prod_acc : 0.0);
    printf("Accepted %ld of %ld steps; kernels launched: %lld\n", nsuc, cfg.lj, (long long)mcb_launch_count(h));
    if (!cfg.result.empty()) {                                                            // result box (:587-611) = next run's input
        FILE *f = fopen(cfg.result.c_str(), "w");
        if (f) {
            fprintf(f, "  %.17g\n  %d\n%20.10E\n%s\n", box.boxl, box.n, toteng, box.id.c_str());
            for (int j = 0; j < box.n; ++j)
                fprintf(f, " %.17g %.17g %.17g %.17g %.17g %.17g\n", box.com[3 * j], box.com[3 * j + 1], box.com[3 * j + 2], box.hoek[3 * j], box.hoek[3 * j + 1], box.hoek[3 * j + 2]);
            fprintf(f, " SOLUTE\n  %d\n%s\n", box.nsol, box.sol_name.c_str());
            for (int j = 0; j < box.nsol; ++j) fprintf(f, " %-3s %.17g %.17g %.17g\n", box.sol_sym[j].c_str(), box.solute[3 * j], box.solute[3 * j + 1], box.solute[3 * j + 2]);
            fprintf(f, "  %.17g\n  %d\n", box.e_sol, (int)box.dih_lines.size());
            for (auto &l : box.dih_lines) fprintf(f, "%s\n", l.c_str());
            fclose(f);
        }
    }
    mcb_destroy(h);
    return 0;
}
