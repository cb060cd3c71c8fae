// parse.cpp -- dense CSV loader with the reference's semantics (reference parse.cpp:10-43), parsed
// with strtol/strtof over a large read buffer instead of a stringstream per line.
#include "parse.hpp"

#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <iostream>
#include <stdexcept>
#include <string>

using namespace std;

namespace {

// getline() over a FILE* with a big buffer; returns false at EOF with nothing read.
struct LineReader {
    FILE* fp;
    std::string buf;
    explicit LineReader(FILE* f) : fp(f) { setvbuf(fp, nullptr, _IOFBF, 1 << 22); }
    bool next(std::string& line)
    {
        line.clear();
        char chunk[1 << 16];
        bool any = false;
        while (fgets(chunk, sizeof(chunk), fp)) {
            any = true;
            size_t len = strlen(chunk);
            if (len && chunk[len - 1] == '\n') { line.append(chunk, len - 1); return true; }
            line.append(chunk, len);
        }
        return any;
    }
};

void parse_line(const std::string& line, int* label, float* row, int num_attributes)
{
    const char* p = line.c_str();
    const char* end = p + line.size();
    // first cell: stoi
    const char* comma = (const char*)memchr(p, ',', (size_t)(end - p));
    {
        char* q = nullptr;
        errno = 0;
        long v = strtol(p, &q, 10);
        if (q == p) throw std::invalid_argument("stoi");
        if (errno == ERANGE || v < -2147483648L || v > 2147483647L) throw std::out_of_range("stoi");
        *label = (int)v;
    }
    if (!comma) return;
    p = comma + 1;
    int k = 0;
    // remaining cells: stof each.  A trailing empty cell after the last comma is not a cell
    // (getline(lineStream, cell, ',') returns false at end of stream).
    while (p <= end) {
        if (p == end) break;
        const char* next = (const char*)memchr(p, ',', (size_t)(end - p));
        char* q = nullptr;
        errno = 0;
        float v = strtof(p, &q);
        if (q == p || (next && q > next)) throw std::invalid_argument("stof");
        if (k >= num_attributes) throw std::out_of_range("populate_data: row has more cells than num_attributes");
        row[k++] = v;
        if (!next) break;
        p = next + 1;
    }
}

void load_rows(std::vector<float>& x, std::vector<int>& y, int first_row, int num_rows, int num_attributes, const char* path)
{
    FILE* fp = fopen(path, "r");
    if (!fp) {
        cout << "Couldn't open file\n";     // parse.cpp:14-18
        exit(-1);
    }
    LineReader rd(fp);
    std::string line;
    for (int i = 0; i < first_row; i++) {
        if (!rd.next(line)) { fclose(fp); throw std::invalid_argument("stoi"); }
    }
    for (int i = 0; i < num_rows; i++) {
        if (!rd.next(line)) line.clear();    // short file: the reference's stoi("") throws
        try {
            parse_line(line, &y[i], &x[(size_t)i * num_attributes], num_attributes);
        } catch (...) {
            fclose(fp);
            throw;
        }
    }
    fclose(fp);
}

}  // namespace

void populate_data(vector<float>& x, vector<int>& y, int num_train_data, int num_attributes, char input_file_name[200])
{
    load_rows(x, y, 0, num_train_data, num_attributes, input_file_name);
}

void populate_data_rows(vector<float>& x, vector<int>& y, int first_row, int num_rows, int num_attributes,
                        const char* input_file_name)
{
    load_rows(x, y, first_row, num_rows, num_attributes, input_file_name);
}
