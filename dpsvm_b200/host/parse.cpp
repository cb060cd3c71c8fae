// parse.cpp -- dense CSV loader with the reference's semantics (reference parse.cpp:10-43):
//   per line: first cell through stoi -> label, every further comma-separated cell through stof into
//   x[row*d + k]; exactly num_train_data lines are consumed; short rows leave the remaining cells as
//   they were; an unopenable file prints "Couldn't open file" and exit(-1); a missing line or an
//   unparsable cell throws std::invalid_argument, an out-of-range one std::out_of_range.
// Built for throughput (the reference spends a stringstream per line and a std::string per cell, on every
// rank, over the whole file): the file is mapped, line starts are found with one memchr pass, and rows
// are parsed in parallel with std::from_chars -- strtof only where its behaviour could differ.
#include "parse.hpp"

#include <errno.h>
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <charconv>
#include <cmath>
#include <exception>
#include <iostream>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

using namespace std;

namespace {

inline bool is_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\v' || c == '\f' || c == '\r'; }

// std::stoi on the cell [p, end): leading blanks, optional sign, decimal digits, rest ignored.
int cell_stoi(const char* p, const char* end)
{
    while (p < end && is_space(*p)) ++p;
    bool neg = false;
    if (p < end && (*p == '+' || *p == '-')) { neg = (*p == '-'); ++p; }
    if (p >= end || *p < '0' || *p > '9') throw std::invalid_argument("stoi");
    long long v = 0;
    while (p < end && *p >= '0' && *p <= '9') {
        v = v * 10 + (*p - '0');
        if (v > 2147483648LL) throw std::out_of_range("stoi");
        ++p;
    }
    if (neg) v = -v;
    if (v > 2147483647LL || v < -2147483648LL) throw std::out_of_range("stoi");
    return (int)v;
}

// std::stof on the cell [p, end).  Fast path: from_chars on a plain decimal that yields a normal value or
// zero; anything else (hex floats, inf/nan spellings, subnormals, overflow, leading '+') goes through
// strtof on a bounded copy so that ERANGE / no-conversion behave exactly as stof does.
float cell_stof(const char* p, const char* end)
{
    while (p < end && is_space(*p)) ++p;
    if (p < end) {
        float v;
        const auto r = std::from_chars(p, end, v, std::chars_format::general);
        if (r.ec == std::errc() && (v == 0.0f || std::isnormal(v))) {
            const char c = *p;
            const bool hexish = (end - p > 2) && ((c == '0' && (p[1] == 'x' || p[1] == 'X')) || ((c == '-') && p[1] == '0' && (p[2] == 'x' || p[2] == 'X')));
            if (!hexish) return v;
        }
    }
    char buf[128];
    const size_t len = std::min((size_t)(end - p), sizeof(buf) - 1);
    memcpy(buf, p, len);
    buf[len] = 0;
    char* q = nullptr;
    errno = 0;
    const float v = strtof(buf, &q);
    if (q == buf) throw std::invalid_argument("stof");
    if (errno == ERANGE) throw std::out_of_range("stof");
    return v;
}

void parse_row(const char* p, const char* end, int* label, float* row, int num_attributes)
{
    if (end > p && end[-1] == '\n') --end;
    if (end < p) end = p;
    const char* comma = (const char*)memchr(p, ',', (size_t)(end - p));
    *label = cell_stoi(p, comma ? comma : end);
    if (!comma) return;
    p = comma + 1;
    int k = 0;
    // getline(lineStream, cell, ',') yields no cell once the stream is at its end, so a trailing comma adds none
    while (p < end) {
        const char* next = (const char*)memchr(p, ',', (size_t)(end - p));
        const char* cell_end = next ? next : end;
        const float v = cell_stof(p, cell_end);
        if (k >= num_attributes) throw std::out_of_range("populate_data: row has more cells than num_attributes");
        row[k++] = v;
        if (!next) break;
        p = next + 1;
    }
}

struct Mapping {
    const char* data = nullptr;
    size_t size = 0;
    bool mapped = false;
    std::vector<char> owned;
    ~Mapping() { if (mapped && data) munmap(const_cast<char*>(data), size); }
};

bool open_input(const char* path, Mapping& m)
{
    const int fd = open(path, O_RDONLY);
    if (fd < 0) return false;
    struct stat st;
    if (fstat(fd, &st) == 0 && S_ISREG(st.st_mode) && st.st_size > 0) {
        void* p = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (p != MAP_FAILED) {
            madvise(p, (size_t)st.st_size, MADV_SEQUENTIAL);
            m.data = (const char*)p; m.size = (size_t)st.st_size; m.mapped = true;
            close(fd);
            return true;
        }
    }
    // pipes, empty files, mmap refused: plain reads
    char chunk[1 << 16];
    ssize_t got;
    while ((got = read(fd, chunk, sizeof(chunk))) > 0) m.owned.insert(m.owned.end(), chunk, chunk + got);
    close(fd);
    m.data = m.owned.data(); m.size = m.owned.size();
    return true;
}

void load_rows(std::vector<float>& x, std::vector<int>& y, int first_row, int num_rows, int num_attributes, const char* path)
{
    Mapping m;
    if (!open_input(path, m)) {
        cout << "Couldn't open file\n";     // parse.cpp:14-18
        exit(-1);
    }
    // line starts of rows [first_row, first_row + num_rows)
    std::vector<const char*> starts((size_t)num_rows + 1, nullptr);
    const char* p = m.data;
    const char* const end = m.data + m.size;
    int line = 0, have = 0;
    while (line < first_row + num_rows && p < end) {
        if (line >= first_row) starts[(size_t)have++] = p;
        const char* nl = (const char*)memchr(p, '\n', (size_t)(end - p));
        p = nl ? nl + 1 : end;
        ++line;
    }
    starts[(size_t)have] = p;
    if (line < first_row) throw std::invalid_argument("stoi");          // the reference's stoi("") on a missing line

    const int threads = std::max(1, std::min<int>({(int)std::thread::hardware_concurrency(), 16, have / 256 + 1}));
    std::vector<std::exception_ptr> errors((size_t)threads);
    std::vector<int> error_row((size_t)threads, -1);
    auto work = [&](int t) {
        const int lo = (int)((long long)have * t / threads), hi = (int)((long long)have * (t + 1) / threads);
        for (int i = lo; i < hi; ++i) {
            try {
                parse_row(starts[(size_t)i], starts[(size_t)i + 1], &y[(size_t)i], &x[(size_t)i * num_attributes], num_attributes);
            } catch (...) {
                errors[(size_t)t] = std::current_exception();
                error_row[(size_t)t] = i;
                return;
            }
        }
    };
    if (threads == 1) {
        work(0);
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) pool.emplace_back(work, t);
        for (auto& th : pool) th.join();
    }
    // the reference stops at the FIRST bad row; threads own ascending row blocks
    for (int t = 0; t < threads; ++t) if (errors[(size_t)t]) std::rethrow_exception(errors[(size_t)t]);
    if (have < num_rows) throw std::invalid_argument("stoi");           // file shorter than num_train_data lines
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

// The inverse of populate_data: "label,f0,...,f{d-1}" per line, each value in the shortest decimal form that
// stof() reads back to the same fp32 bits (std::to_chars).  The reference has no writer in C++ -- its data
// files come out of scripts/convert_mnist_to_odd_even.py and convert_adult.py in this layout; this is what
// hands the SAME synthetic rows to the reference's own programs (bench.py, tools/refgpu_compare.py).
// Rows are formatted in parallel into per-thread buffers, then written in order.  Returns 0, or -1 on an I/O error.
extern "C" int dpsvm_host_write_csv(const char* path, const float* x, const int* y, int n, int d)
{
    if (!path || !x || !y || n < 0 || d < 0) return -1;
    FILE* fp = fopen(path, "wb");
    if (!fp) return -1;
    const int threads = std::max(1, std::min<int>({(int)std::thread::hardware_concurrency(), 16, n / 256 + 1}));
    const int block = 4096;                       // rows per round and thread
    std::vector<std::string> bufs((size_t)threads);
    int rc = 0;
    for (int base = 0; base < n && rc == 0; base += block * threads) {
        auto work = [&](int t) {
            std::string& b = bufs[(size_t)t];
            b.clear();
            const int lo = std::min(n, base + t * block), hi = std::min(n, lo + block);
            char tmp[64];
            for (int i = lo; i < hi; ++i) {
                auto r = std::to_chars(tmp, tmp + sizeof(tmp), y[i]);
                b.append(tmp, r.ptr);
                const float* row = x + (size_t)i * d;
                for (int k = 0; k < d; ++k) {
                    b.push_back(',');
                    if (row[k] == 0.0f && !std::signbit(row[k])) { b.push_back('0'); continue; }
                    r = std::to_chars(tmp, tmp + sizeof(tmp), row[k]);
                    b.append(tmp, r.ptr);
                }
                b.push_back('\n');
            }
        };
        if (threads == 1) work(0);
        else {
            std::vector<std::thread> pool;
            for (int t = 0; t < threads; ++t) pool.emplace_back(work, t);
            for (auto& th : pool) th.join();
        }
        for (int t = 0; t < threads && rc == 0; ++t)
            if (!bufs[(size_t)t].empty() && fwrite(bufs[(size_t)t].data(), 1, bufs[(size_t)t].size(), fp) != bufs[(size_t)t].size()) rc = -1;
    }
    if (fclose(fp) != 0) rc = -1;
    return rc;
}
