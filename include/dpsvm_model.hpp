// dpsvm_model.hpp -- model file I/O in the reference's two text formats.
#ifndef DPSVM_MODEL_HPP
#define DPSVM_MODEL_HPP

#include <string>
#include <vector>

struct dpsvm_model {
    float gamma = 0.f;
    float b = 0.f;
    bool has_b = false;
    int num_attributes = 0;
    std::vector<float> alpha;   // per support vector
    std::vector<int> y;
    std::vector<float> x;       // [nsv x num_attributes]
};

// reference svmTrainMain.cpp:386-416: "gamma\n b\n" then "alpha,y,x0,...,x{d-1}\n" for every row with
// alpha != 0, default ostream formatting (6 significant digits).  with_b = false writes seq.cpp's
// variant (seq.cpp:295-321: no b line), which is what seq_test.cpp reads (seq_test.cpp:212-249).
// Returns false if the file cannot be opened.
bool dpsvm_write_model(const char* path, const float* x, const int* y, const float* alpha, float b, float gamma,
                       int num_train_data, int num_attributes, bool with_b);

// Reader for both variants (the reference's svmTest source is missing from its checkout; the
// format is defined by the writers above and by seq_test.cpp:212-270).
bool dpsvm_read_model(const char* path, int num_attributes, bool with_b, dpsvm_model* out);

#endif
