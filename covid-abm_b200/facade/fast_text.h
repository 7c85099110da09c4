// covid-abm_b200/facade/fast_text.h -- whitespace-separated text tables read from one in-memory copy of the file.
//
// The reference parses its five population files with getline + istringstream per row (Community.cpp:174-524); at 21.5 M
// people that is minutes. Same grammar here -- rows of blank-separated fields, a row whose leading fields do not parse as
// the expected numbers is a header and is skipped -- but the file is read with one fread and fields are converted in place
// with std::from_chars (no locale, no stream state, no per-row allocation).
#ifndef CABM_FAST_TEXT_H
#define CABM_FAST_TEXT_H
#include <charconv>
#include <cstdio>
#include <string>
#include <string_view>
#include <vector>

namespace cabm_text {

class Table {
public:
    explicit Table(const std::string& path) {
        FILE* f = std::fopen(path.c_str(), "rb");
        if (!f) return;
        std::fseek(f, 0, SEEK_END); const long n = std::ftell(f); std::fseek(f, 0, SEEK_SET);
        buf_.resize(n > 0 ? (size_t) n : 0);
        ok_ = buf_.empty() or std::fread(buf_.data(), 1, buf_.size(), f) == buf_.size();
        std::fclose(f);
        line_end_ = cur_ = buf_.data(); end_ = buf_.data() + buf_.size();
    }
    bool ok() const { return ok_; }

    // advance to the next row; false at end of file
    bool next_row() {
        if (line_end_ >= end_) return false;
        cur_ = line_end_;
        const char* p = cur_;
        while (p < end_ and *p != '\n') ++p;
        row_end_ = p;
        line_end_ = p < end_ ? p + 1 : p;
        return true;
    }
    // next blank-separated field of the row, empty when the row is exhausted
    std::string_view field() {
        while (cur_ < row_end_ and (*cur_ == ' ' or *cur_ == '\t' or *cur_ == '\r')) ++cur_;
        const char* b = cur_;
        while (cur_ < row_end_ and *cur_ != ' ' and *cur_ != '\t' and *cur_ != '\r') ++cur_;
        return std::string_view(b, (size_t) (cur_ - b));
    }
    template <typename T> bool number(T& out) {
        const std::string_view s = field();
        if (s.empty()) return false;
        const char* b = s.data(); const char* e = b + s.size();
        if (*b == '+') ++b;
        const auto r = std::from_chars(b, e, out);
        return r.ec == std::errc() and r.ptr == e;
    }

private:
    std::vector<char> buf_;
    const char* cur_ = nullptr; const char* row_end_ = nullptr; const char* line_end_ = nullptr; const char* end_ = nullptr;
    bool ok_ = false;
};

} // namespace cabm_text
#endif
