/* include/cabm_popfile.h -- the binary population file of cabm_population_save / cabm_load_population_file (cabm.h).
 *
 * Host-only C++, shared by the CUDA library and the test oracle. Layout (little-endian):
 *   header  : "CABMPOP1", u32 n_people, n_loc, n_patron, reserved, u64 n_nbr, n_roster, file_bytes
 *   arrays  : loc_type[L] u8, loc_essential[L] u8, loc_pubrisk[L] u8, loc_risk[L] f32, loc_hospital[L] i32, nbr_off[L+1] i64,
 *             nbr[n_nbr] i32, roster_off[L+1] i64, roster[n_roster] i32, home[N] i32, day[N] i32, age[N] u8, sex[N] u8,
 *             comorbid[N] u8, ltc[N] u8, patron[N * n_patron] i32 -- each padded to a multiple of 8 bytes
 * The load-time immunity attributes are not part of the file: they depend on the run's seed (Community.cpp:212-215). */
#ifndef CABM_POPFILE_H
#define CABM_POPFILE_H
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstdio>
#include <cstring>
#include <string>

#include "cabm.h"

namespace cabm_popfile {

struct Header {
    char magic[8];
    uint32_t n_people, n_loc, n_patron, reserved;
    uint64_t n_nbr, n_roster;
    uint64_t file_bytes;
};
enum { N_ARRAYS = 16 };
inline size_t pad8(size_t n) { return (n + 7) & ~(size_t) 7; }
inline void array_bytes(size_t N, size_t L, size_t n_patron, uint64_t n_nbr, uint64_t n_roster, size_t out[N_ARRAYS]) {
    const size_t b[N_ARRAYS] = { L, L, L, L * 4, L * 4, (L + 1) * 8, (size_t) n_nbr * 4, (L + 1) * 8, (size_t) n_roster * 4,
                                 N * 4, N * 4, N, N, N, N, N * n_patron * 4 };
    for (int i = 0; i < N_ARRAYS; ++i) out[i] = b[i];
}

inline int save(const cabm_population* pop, const char* path) {
    if (!pop || !path || !pop->nbr_off || !pop->roster_off) return CABM_E_ARG;
    if (!pop->loc_type || !pop->loc_essential || !pop->loc_pubrisk || !pop->loc_risk || !pop->loc_hospital || !pop->home || !pop->day || !pop->age ||
        !pop->sex || !pop->comorbid || !pop->ltc || (pop->n_patron && !pop->patron)) return CABM_E_ARG;
    const int64_t n_nbr = pop->nbr_off[pop->n_loc], n_roster = pop->roster_off[pop->n_loc];
    if (n_nbr < 0 || n_roster < 0 || (n_nbr && !pop->nbr) || (n_roster && !pop->roster)) return CABM_E_ARG;
    Header hd; std::memset(&hd, 0, sizeof(hd));
    std::memcpy(hd.magic, "CABMPOP1", 8);
    hd.n_people = pop->n_people; hd.n_loc = pop->n_loc; hd.n_patron = pop->n_patron; hd.n_nbr = (uint64_t) n_nbr; hd.n_roster = (uint64_t) n_roster;
    size_t bytes[N_ARRAYS]; array_bytes(pop->n_people, pop->n_loc, pop->n_patron, hd.n_nbr, hd.n_roster, bytes);
    const void* ptr[N_ARRAYS] = { pop->loc_type, pop->loc_essential, pop->loc_pubrisk, pop->loc_risk, pop->loc_hospital, pop->nbr_off, pop->nbr,
                                  pop->roster_off, pop->roster, pop->home, pop->day, pop->age, pop->sex, pop->comorbid, pop->ltc, pop->patron };
    size_t total = sizeof(hd);
    for (int i = 0; i < N_ARRAYS; ++i) total += pad8(bytes[i]);
    hd.file_bytes = total;
    const std::string tmp = std::string(path) + ".tmp" + std::to_string((long) getpid());       /* written under a temporary name, then renamed */
    FILE* f = std::fopen(tmp.c_str(), "wb");
    if (!f) return CABM_E_ARG;
    bool ok = std::fwrite(&hd, sizeof(hd), 1, f) == 1;
    static const char zeros[8] = { 0 };
    for (int i = 0; i < N_ARRAYS && ok; ++i) {
        if (bytes[i]) ok = std::fwrite(ptr[i], 1, bytes[i], f) == bytes[i];
        const size_t pad = pad8(bytes[i]) - bytes[i];
        if (ok && pad) ok = std::fwrite(zeros, 1, pad, f) == pad;
    }
    ok = (std::fclose(f) == 0) && ok;
    if (!ok || std::rename(tmp.c_str(), path) != 0) { std::remove(tmp.c_str()); return CABM_E_ARG; }
    return CABM_OK;
}

/* a population file mapped read-only with a cabm_population laid over it (the immunity attributes stay NULL: the caller sets them) */
struct Mapped {
    void* map = nullptr; size_t bytes = 0;
    cabm_population pop;
    Mapped() { std::memset(&pop, 0, sizeof(pop)); }
    Mapped(const Mapped&) = delete;
    Mapped& operator=(const Mapped&) = delete;
    ~Mapped() { if (map) munmap(map, bytes); }
    int open_file(const char* path, std::string* err) {
        const int fd = open(path, O_RDONLY);
        if (fd < 0) { *err = std::string("population file not found: ") + path; return CABM_E_ARG; }
        struct stat st;
        if (fstat(fd, &st) != 0 || (size_t) st.st_size < sizeof(Header)) { close(fd); *err = "population file: too short for its header"; return CABM_E_ARG; }
        void* m = mmap(nullptr, (size_t) st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
        close(fd);
        if (m == MAP_FAILED) { *err = "population file: mmap failed"; return CABM_E_NOMEM; }
        map = m; bytes = (size_t) st.st_size;
        const Header& hd = *static_cast<const Header*>(map);
        if (std::memcmp(hd.magic, "CABMPOP1", 8) != 0) { *err = "population file: bad magic (not a CABMPOP1 file)"; return CABM_E_ARG; }
        if (hd.file_bytes != (uint64_t) bytes) { *err = "population file: size does not match its header (truncated?)"; return CABM_E_ARG; }
        size_t sz[N_ARRAYS]; array_bytes(hd.n_people, hd.n_loc, hd.n_patron, hd.n_nbr, hd.n_roster, sz);
        const unsigned char* base = static_cast<const unsigned char*>(map);
        const void* ptr[N_ARRAYS];
        size_t off = sizeof(Header);
        for (int i = 0; i < N_ARRAYS; ++i) { ptr[i] = base + off; off += pad8(sz[i]); if (off > bytes) break; }
        if (off != bytes) { *err = "population file: array sizes do not add up to the file size"; return CABM_E_ARG; }
        pop.n_people = hd.n_people; pop.n_loc = hd.n_loc; pop.n_patron = hd.n_patron;
        pop.loc_type = (const uint8_t*) ptr[0]; pop.loc_essential = (const uint8_t*) ptr[1]; pop.loc_pubrisk = (const uint8_t*) ptr[2];
        pop.loc_risk = (const float*) ptr[3]; pop.loc_hospital = (const int32_t*) ptr[4];
        pop.nbr_off = (const int64_t*) ptr[5]; pop.nbr = (const int32_t*) ptr[6]; pop.roster_off = (const int64_t*) ptr[7]; pop.roster = (const int32_t*) ptr[8];
        pop.home = (const int32_t*) ptr[9]; pop.day = (const int32_t*) ptr[10]; pop.age = (const uint8_t*) ptr[11]; pop.sex = (const uint8_t*) ptr[12];
        pop.comorbid = (const uint8_t*) ptr[13]; pop.ltc = (const uint8_t*) ptr[14]; pop.patron = (const int32_t*) ptr[15];
        if (pop.nbr_off[hd.n_loc] != (int64_t) hd.n_nbr || pop.roster_off[hd.n_loc] != (int64_t) hd.n_roster) { *err = "population file: CSR offsets do not match the header"; return CABM_E_ARG; }
        return CABM_OK;
    }
};

} /* namespace cabm_popfile */
#endif
