// CPU-side checks of the host layer (no GPU): populate_data semantics (reference parse.cpp:10-43) and the
// model file formats (reference svmTrainMain.cpp:386-416, seq.cpp:295-321).  Driven by tests/test_host_cpp.py.
//   test_host parse <csv> <n> <d>                 -> prints y and x as %d / %.9g, one row per line
//   test_host parse_rows <csv> <first> <n> <d>
//   test_host write_model <csv> <n> <d> <alpha.bin> <b> <gamma> <with_b> <out>
//   test_host read_model <model> <d> <with_b>     -> prints gamma, b, nsv, then alpha,y,x...
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <stdexcept>
#include <string>
#include <vector>

#include "dpsvm_model.hpp"
#include "parse.hpp"
#include "svmTrainMain.hpp"

state_model state;      // the driver defines it (reference svmTrainMain.cpp:18); libdpsvm_host.so reads it

static void dump(const std::vector<float>& x, const std::vector<int>& y, int n, int d)
{
    for (int i = 0; i < n; i++) {
        printf("%d", y[i]);
        for (int k = 0; k < d; k++) printf(",%.9g", x[(size_t)i * d + k]);
        printf("\n");
    }
}

int main(int argc, char** argv)
{
    if (argc < 2) return 2;
    const std::string cmd = argv[1];
    try {
        if (cmd == "parse" && argc == 5) {
            const int n = atoi(argv[3]), d = atoi(argv[4]);
            std::vector<float> x((size_t)n * d, 0);
            std::vector<int> y(n, 0);
            char name[200];
            snprintf(name, sizeof(name), "%s", argv[2]);
            populate_data(x, y, n, d, name);
            dump(x, y, n, d);
            return 0;
        }
        if (cmd == "parse_rows" && argc == 6) {
            const int first = atoi(argv[3]), n = atoi(argv[4]), d = atoi(argv[5]);
            std::vector<float> x((size_t)n * d, 0);
            std::vector<int> y(n, 0);
            populate_data_rows(x, y, first, n, d, argv[2]);
            dump(x, y, n, d);
            return 0;
        }
        if (cmd == "write_model" && argc == 10) {
            const int n = atoi(argv[3]), d = atoi(argv[4]);
            std::vector<float> x((size_t)n * d, 0), alpha(n);
            std::vector<int> y(n, 0);
            char name[200];
            snprintf(name, sizeof(name), "%s", argv[2]);
            populate_data(x, y, n, d, name);
            FILE* fp = fopen(argv[5], "rb");
            if (!fp || fread(alpha.data(), sizeof(float), n, fp) != (size_t)n) return 3;
            fclose(fp);
            const bool ok = dpsvm_write_model(argv[9], x.data(), y.data(), alpha.data(), (float)atof(argv[6]), (float)atof(argv[7]),
                                              n, d, atoi(argv[8]) != 0);
            return ok ? 0 : 4;
        }
        if (cmd == "read_model" && argc == 5) {
            dpsvm_model m;
            if (!dpsvm_read_model(argv[2], atoi(argv[3]), atoi(argv[4]) != 0, &m)) return 5;
            printf("%.9g %.9g %zu\n", m.gamma, m.b, m.alpha.size());
            for (size_t i = 0; i < m.alpha.size(); i++) {
                printf("%.9g,%d", m.alpha[i], m.y[i]);
                for (int k = 0; k < m.num_attributes; k++) printf(",%.9g", m.x[i * m.num_attributes + k]);
                printf("\n");
            }
            return 0;
        }
    } catch (const std::invalid_argument& e) {
        printf("invalid_argument:%s\n", e.what());
        return 10;
    } catch (const std::out_of_range& e) {
        printf("out_of_range:%s\n", e.what());
        return 11;
    }
    return 2;
}
