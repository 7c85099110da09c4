// covid-abm_b200/facade/pop_cache.h -- the five population text files as flat arrays, and a binary cache of those arrays.
//
// "fast loaders for the 5 text files, binary cache of the CSR build". The loaders of the facade
// (Community::loadLocations / loadPopulation, the reference's Community.cpp:174-524) work in two steps: the text is parsed into
// the row arrays below (same grammar and the same error messages as the reference's parsers, fast_text.h), then the proxies are
// built from the arrays in row order, drawing from RNG exactly where the reference does. With CABM_POP_CACHE=<directory> the
// arrays of a file set are written to <directory>/cabm_{loc,pop}_<key>.bin after the first parse and read back instead of the
// text afterwards; <key> hashes the files' paths, sizes and modification times, so an edited file is parsed again. The cache
// holds what the FILES say, nothing that depends on the run's seed: the random household riskiness and the per-person immunity
// attributes are drawn at build time as before, in the same order.
#ifndef CABM_POP_CACHE_H
#define CABM_POP_CACHE_H
#include <sys/stat.h>
#include <unistd.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "fast_text.h"

namespace cabm_pop {

// one row of the locations file (Community.cpp:385-470): id x y type essential hfid [compliance [public]]
struct LocRows {
    std::vector<double> x, y, compliance;            // compliance: meaningful where flags bit 0 is set
    std::vector<int32_t> hfid;
    std::vector<uint8_t> type;                       // LocationType as the facade's enum value
    std::vector<uint8_t> flags;                      // bit 0: a compliance number followed; bit 1: essential; bits 2-3: public 0 none, 1 low, 2 high
    std::vector<int32_t> net;                        // the network file: pairs (a, b)
};
// population, comorbidity and public-activity files (Community.cpp:196-300)
struct PopRows {
    std::vector<int32_t> hid, did, sex, age;
    std::vector<int32_t> comorbid;                   // person ids with comorbidity == 1, in file order
    std::vector<int32_t> pub_pid;                    // one entry per public-activity row
    std::vector<int64_t> pub_off;                    // [rows + 1] into pub_loc
    std::vector<int32_t> pub_loc;
};

enum { LT_HOUSE_ = 0, LT_WORK_, LT_SCHOOL_, LT_HOSPITAL_, LT_NURSINGHOME_, LT_UNKNOWN_ };

// ---- text -> rows: the reference's grammar and error texts ----
inline bool parse_locations(const std::string& locationFilename, const std::string& networkFilename, LocRows& R, const int type_codes[5]) {
    using std::cerr; using std::endl;
    cabm_text::Table in(locationFilename);
    if (!in.ok()) { cerr << "ERROR: locations file " << locationFilename << " not found." << endl; return false; }
    while (in.next_row()) {
        int locid, hfid; double x, y, compliance = 0.0;
        if (!(in.number(locid) and in.number(x) and in.number(y))) continue;               // header or malformed line
        const std::string_view type_s = in.field(), ess_s = in.field();
        if (type_s.empty() or ess_s.empty() or !in.number(hfid)) continue;
        if (locid != (int) R.x.size()) { cerr << "ERROR: Location ID's must be sequential integers starting at 0" << endl; return false; }
        const int lt = type_s == "h" ? LT_HOUSE_ : type_s == "w" ? LT_WORK_ : type_s == "s" ? LT_SCHOOL_ : type_s == "hf" ? LT_HOSPITAL_ : type_s == "n" ? LT_NURSINGHOME_ : LT_UNKNOWN_;
        if (lt == LT_UNKNOWN_) { cerr << "ERROR: Parsed unknown location type: " << type_s << " from location file: " << locationFilename << endl; return false; }
        const bool known_ess = ess_s == "y" or ess_s == "NA" or ess_s == "n";
        if (!known_ess) { cerr << "ERROR: Unknown value for \"essential\" status: " << ess_s << " from location file: " << locationFilename << endl; return false; }
        const bool have_compliance = in.number(compliance);                                 // a field that is not a number ends the row, as a failed stream does
        unsigned int pub = 0;
        const std::string_view pub_s = have_compliance ? in.field() : std::string_view();
        if (not pub_s.empty() and pub_s != "N") {
            pub = pub_s == "H" ? 2u : pub_s == "L" ? 1u : 3u;
            if (pub == 3u) { cerr << "ERROR: Unknown value for \"public transmission\" status: " << pub_s << endl; return false; }
        }
        R.x.push_back(x); R.y.push_back(y); R.compliance.push_back(have_compliance ? compliance : 0.0); R.hfid.push_back(hfid);
        R.type.push_back((uint8_t) type_codes[lt]);
        R.flags.push_back((uint8_t) ((have_compliance ? 1u : 0u) | (ess_s != "n" ? 2u : 0u) | (pub << 2)));
    }
    cabm_text::Table net(networkFilename);
    if (!net.ok()) { cerr << "ERROR: network file " << networkFilename << " not found." << endl; return false; }
    while (net.next_row()) {
        int a, b;
        if (net.number(a) and net.number(b)) { R.net.push_back(a); R.net.push_back(b); }
    }
    return true;
}

inline bool parse_population(const std::string& populationFilename, const std::string& comorbidityFilename, const std::string& publicActivityFilename, PopRows& R) {
    using std::cerr; using std::endl;
    {
        cabm_text::Table in(populationFilename);
        if (!in.ok()) { cerr << "ERROR: " << populationFilename << " not found." << endl; return false; }
        while (in.next_row()) {
            int pid, hid, sex, age, did;
            if (!(in.number(pid) and in.number(hid) and in.number(sex) and in.number(age) and in.number(did))) continue;
            if (pid != (int) R.hid.size()) { cerr << "ERROR: Person ID's must be sequential integers starting at 0" << endl; return false; }
            R.hid.push_back(hid); R.did.push_back(did); R.sex.push_back(sex); R.age.push_back(age);
        }
    }
    if (comorbidityFilename.length() > 0) {
        cabm_text::Table in(comorbidityFilename);
        if (!in.ok()) { cerr << "ERROR: " << comorbidityFilename << " not found." << endl; return false; }
        while (in.next_row()) {
            int pid, sex, age, com;
            if (in.number(pid) and in.number(sex) and in.number(age) and in.number(com) and com == 1) R.comorbid.push_back(pid);
        }
    }
    R.pub_off.assign(1, 0);
    if (publicActivityFilename.length() > 0) {
        cabm_text::Table in(publicActivityFilename);
        if (!in.ok()) { cerr << "ERROR: " << publicActivityFilename << " not found." << endl; return false; }
        while (in.next_row()) {
            int pid, locid;
            if (!in.number(pid)) continue;
            R.pub_pid.push_back(pid);
            while (in.number(locid)) R.pub_loc.push_back(locid);
            R.pub_off.push_back((int64_t) R.pub_loc.size());
        }
    }
    return true;
}

// ---- rows <-> binary cache ----
inline uint64_t fnv1a(uint64_t h, const void* p, size_t n) {
    const unsigned char* b = (const unsigned char*) p;
    for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}
// key of a file set: every path with its size and modification time (an absent optional file contributes its empty name)
inline bool file_set_key(const std::vector<std::string>& paths, uint64_t& key) {
    key = 1469598103934665603ull;
    for (const std::string& p : paths) {
        key = fnv1a(key, p.data(), p.size() + 1);
        if (p.empty()) continue;
        struct stat st;
        if (stat(p.c_str(), &st) != 0) return false;
        const int64_t sz = (int64_t) st.st_size, mt = (int64_t) st.st_mtime, ns = (int64_t) st.st_mtim.tv_nsec;
        key = fnv1a(key, &sz, sizeof(sz)); key = fnv1a(key, &mt, sizeof(mt)); key = fnv1a(key, &ns, sizeof(ns));
    }
    return true;
}
inline std::string cache_path(const char* kind, uint64_t key) {
    const char* dir = std::getenv("CABM_POP_CACHE");
    if (!dir or !*dir) return std::string();
    char name[64]; std::snprintf(name, sizeof(name), "/cabm_%s_%016llx.bin", kind, (unsigned long long) key);
    return std::string(dir) + name;
}

struct Writer {
    FILE* f; bool ok;
    explicit Writer(const std::string& path) : f(std::fopen(path.c_str(), "wb")), ok(f != nullptr) {}
    ~Writer() { if (f) std::fclose(f); }
    template <typename T> void vec(const std::vector<T>& v) {
        const uint64_t n = v.size();
        ok = ok and std::fwrite(&n, sizeof(n), 1, f) == 1 and (n == 0 or std::fwrite(v.data(), sizeof(T), n, f) == n);
    }
    void u64(uint64_t x) { ok = ok and std::fwrite(&x, sizeof(x), 1, f) == 1; }
};
struct Reader {
    FILE* f; bool ok;
    explicit Reader(const std::string& path) : f(std::fopen(path.c_str(), "rb")), ok(f != nullptr) {}
    ~Reader() { if (f) std::fclose(f); }
    template <typename T> void vec(std::vector<T>& v) {
        uint64_t n = 0;
        ok = ok and std::fread(&n, sizeof(n), 1, f) == 1 and n < (1ull << 36);
        if (!ok) return;
        v.resize(n);
        ok = n == 0 or std::fread(v.data(), sizeof(T), n, f) == n;
    }
    uint64_t u64() { uint64_t x = 0; ok = ok and std::fread(&x, sizeof(x), 1, f) == 1; return x; }
};
static const uint64_t CACHE_MAGIC = 0x3143504d42414321ull;      // "!CABMPC1" little-endian: bump the last byte when the row layout changes

// a cache file is written under a temporary name and renamed, so a reader never sees a half-written one
template <typename Rows, typename Fn> bool save_rows(const std::string& path, uint64_t key, const Rows& R, Fn each) {
    if (path.empty()) return false;
    const std::string tmp = path + ".tmp" + std::to_string((long) getpid());
    {
        Writer w(tmp);
        if (!w.ok) return false;
        w.u64(CACHE_MAGIC); w.u64(key);
        each(w, const_cast<Rows&>(R));
        w.u64(CACHE_MAGIC);
        if (!w.ok) { std::remove(tmp.c_str()); return false; }
    }
    if (std::rename(tmp.c_str(), path.c_str()) != 0) { std::remove(tmp.c_str()); return false; }
    return true;
}
template <typename Rows, typename Fn> bool load_rows(const std::string& path, uint64_t key, Rows& R, Fn each) {
    if (path.empty()) return false;
    Reader r(path);
    if (!r.ok or r.u64() != CACHE_MAGIC or r.u64() != key or !r.ok) return false;
    each(r, R);
    return r.ok and r.u64() == CACHE_MAGIC and r.ok;
}
template <typename IO> void each_loc(IO& io, LocRows& R) { io.vec(R.x); io.vec(R.y); io.vec(R.compliance); io.vec(R.hfid); io.vec(R.type); io.vec(R.flags); io.vec(R.net); }
template <typename IO> void each_pop(IO& io, PopRows& R) { io.vec(R.hid); io.vec(R.did); io.vec(R.sex); io.vec(R.age); io.vec(R.comorbid); io.vec(R.pub_pid); io.vec(R.pub_off); io.vec(R.pub_loc); }

// rows of a file set: from the cache when CABM_POP_CACHE names a directory that holds a current one, else from the text (and
// then into the cache). *from_cache tells which.
inline bool location_rows(const std::string& loc, const std::string& net, const int type_codes[5], LocRows& R, bool* from_cache) {
    uint64_t key = 0; *from_cache = false;
    const bool keyed = file_set_key({ loc, net }, key);
    key = fnv1a(key, type_codes, 5 * sizeof(int));
    const std::string path = keyed ? cache_path("loc", key) : std::string();
    if (load_rows(path, key, R, [](Reader& r, LocRows& x) { each_loc(r, x); }) and R.y.size() == R.x.size() and R.compliance.size() == R.x.size() and
        R.hfid.size() == R.x.size() and R.type.size() == R.x.size() and R.flags.size() == R.x.size() and R.net.size() % 2 == 0) { *from_cache = true; return true; }
    R = LocRows();
    if (!parse_locations(loc, net, R, type_codes)) return false;
    save_rows(path, key, R, [](Writer& w, LocRows& x) { each_loc(w, x); });
    return true;
}
inline bool population_rows(const std::string& pop, const std::string& com, const std::string& pub, PopRows& R, bool* from_cache) {
    uint64_t key = 0; *from_cache = false;
    const bool keyed = file_set_key({ pop, com, pub }, key);
    const std::string path = keyed ? cache_path("pop", key) : std::string();
    if (load_rows(path, key, R, [](Reader& r, PopRows& x) { each_pop(r, x); }) and R.did.size() == R.hid.size() and R.sex.size() == R.hid.size() and
        R.age.size() == R.hid.size() and R.pub_off.size() == R.pub_pid.size() + 1 and (R.pub_off.empty() or R.pub_off.back() == (int64_t) R.pub_loc.size())) { *from_cache = true; return true; }
    R = PopRows();
    if (!parse_population(pop, com, pub, R)) return false;
    save_rows(path, key, R, [](Writer& w, PopRows& x) { each_pop(w, x); });
    return true;
}

} // namespace cabm_pop
#endif
