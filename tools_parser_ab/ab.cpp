// TOOL: driver of the A/B parser timing, see tools_parser_ab/run.sh.
#include <algorithm>
#include <cstdio>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>
namespace tk { void* run_make_new(const std::string&); double run_new(void*, const std::string&, unsigned long long*); }
namespace tk_old { void* run_make_old(const std::string&); double run_old(void*, const std::string&, unsigned long long*); }
static std::string slurp(const char* p) { std::ifstream f(p, std::ios::binary); std::stringstream ss; ss << f.rdbuf(); return ss.str(); }
int main(int argc, char** argv) {
    std::string fa = slurp(argv[1]), mdf = slurp(argv[2]);
    int reps = argc > 3 ? atoi(argv[3]) : 30;
    void* a = tk::run_make_new(fa); void* b = tk_old::run_make_old(fa);
    unsigned long long ha = 0, hb = 0;
    tk::run_new(a, mdf, &ha); tk_old::run_old(b, mdf, &hb);
    printf("checksum new %llx old %llx %s\n", ha, hb, ha == hb ? "SAME" : "DIFFERENT");
    std::vector<double> tn, to, ratio;
    for (int r = 0; r < reps; r++) { double x = tk::run_new(a, mdf, nullptr), y = tk_old::run_old(b, mdf, nullptr); tn.push_back(x); to.push_back(y); ratio.push_back(x / y); }
    std::sort(tn.begin(), tn.end()); std::sort(to.begin(), to.end()); std::sort(ratio.begin(), ratio.end());
    printf("new min %.1f med %.1f | old min %.1f med %.1f | paired ratio med %.3f\n", tn[0], tn[reps / 2], to[0], to[reps / 2], ratio[reps / 2]);
}
