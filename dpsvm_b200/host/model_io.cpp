// model_io.cpp -- model files in the reference's text formats (include/dpsvm_model.hpp).
#include "dpsvm_model.hpp"

#include <fstream>
#include <sstream>

using namespace std;

bool dpsvm_write_model(const char* path, const float* x, const int* y, const float* alpha, float b, float gamma,
                       int num_train_data, int num_attributes, bool with_b)
{
    ofstream model_file;
    model_file.open(path);
    if (!model_file.is_open()) return false;
    model_file << gamma << "\n";                       // svmTrainMain.cpp:393 / seq.cpp:302
    if (with_b) model_file << b << "\n";               // svmTrainMain.cpp:394 (seq.cpp has no b line)
    for (int i = 0; i < num_train_data; i++) {
        if (alpha[i] != 0) {                           // svmTrainMain.cpp:397
            model_file << alpha[i] << "," << y[i];
            const size_t idx = (size_t)i * num_attributes;
            for (int j = 0; j < num_attributes; j++) model_file << "," << x[idx + j];
            model_file << "\n";
        }
    }
    model_file.close();
    return true;
}

bool dpsvm_read_model(const char* path, int num_attributes, bool with_b, dpsvm_model* out)
{
    ifstream file(path);
    if (!file.is_open()) return false;
    string line;
    *out = dpsvm_model();
    out->num_attributes = num_attributes;
    if (!getline(file, line)) return false;
    out->gamma = stof(line);                           // seq_test.cpp:224-226
    if (with_b) {
        if (!getline(file, line)) return false;
        out->b = stof(line);
        out->has_b = true;
    }
    while (getline(file, line)) {
        if (line.empty()) continue;
        stringstream ls(line);
        string cell;
        getline(ls, cell, ',');
        out->alpha.push_back(stof(cell));              // seq_test.cpp:235-236
        getline(ls, cell, ',');
        out->y.push_back(stoi(cell));                  // :238-239
        const size_t base = out->x.size();
        out->x.resize(base + (size_t)num_attributes, 0.0f);
        int k = 0;
        while (getline(ls, cell, ',')) {
            if (k < num_attributes) out->x[base + k] = stof(cell);
            k++;
        }
    }
    return true;
}
