// parse.hpp -- dense CSV loader, same signature as the reference (reference parse.hpp:6).
#ifndef PARSE
#define PARSE

#include <vector>

// Reads exactly num_train_data lines of "label,f0,...,f{d-1}" into y[i] and x[i*d + k]
// (reference parse.cpp:10-43).  Short rows leave the remaining (pre-zeroed) cells untouched.
// Unopenable file: prints "Couldn't open file" and exit(-1), as the reference does.
// Malformed / missing cells throw std::invalid_argument like the reference's stoi/stof.
// Unlike the reference, a row with more than num_attributes features is an error rather than a
// write past the end of the row.
extern void populate_data(std::vector<float> &x, std::vector<int> &y, int num_train_data, int num_attributes,char input_file_name[200]) ;

// Same, restricted to rows [first_row, first_row + num_rows): x and y hold only those rows.
// Lets each GPU rank parse its own shard instead of the whole file (the reference parses the full
// file on every rank, svmTrainMain.cpp:180).
extern void populate_data_rows(std::vector<float> &x, std::vector<int> &y, int first_row, int num_rows,
                               int num_attributes, const char* input_file_name);

// The inverse: writes n rows in the layout populate_data reads (the layout the reference's converters produce,
// scripts/convert_mnist_to_odd_even.py:24-26, convert_adult.py:24-31); values round-trip bit for bit.  0 or -1.
extern "C" int dpsvm_host_write_csv(const char* path, const float* x, const int* y, int n, int d);

#endif
