// covid-abm_b200/facade/Community.cpp -- bodies of the drop-in Community / Person / Location proxies.
//
// Replaces the reference's Community.cpp, Person.cpp and Location.cpp in a driver build: the loaders read the
// reference's own file formats (Community.cpp:174-524), everything that happens inside a simulated day is one C-ABI
// call into the CUDA library (include/cabm.h). CABM_FN(x) prepends the library prefix, so the same file links against
// the product (cabm_) and -- for CPU-only tests of this host logic -- against the oracle (cabm_oracle_).
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <gsl/gsl_rng.h>
#include <gsl/gsl_randist.h>
#include "fast_text.h"
#include "pop_cache.h"
#include "Community.h"
#include "Person.h"
#include "../../include/cabm.h"

#ifndef CABM_PREFIX
#define CABM_PREFIX cabm_
#endif
#define CABM_CAT2(a, b) a##b
#define CABM_CAT(a, b) CABM_CAT2(a, b)
#define CABM_FN(name) CABM_CAT(CABM_PREFIX, name)
extern "C" {
#define CABM_DECL(ret, name, ...) ret CABM_FN(name)(__VA_ARGS__)
CABM_DECL(int, create, cabm_t**); CABM_DECL(int, destroy, cabm_t*); CABM_DECL(const char*, last_error, const cabm_t*);
CABM_DECL(int, load_population, cabm_t*, const cabm_population*); CABM_DECL(int, set_params, cabm_t*, const cabm_params*);
CABM_DECL(int, set_timed_intervention, cabm_t*, int, const double*, size_t);
CABM_DECL(int, tick, cabm_t*, int, int, int); CABM_DECL(int, infect, cabm_t*, uint32_t, int, int, int32_t*);
CABM_DECL(int, infect_batch, cabm_t*, const uint32_t*, const uint8_t*, size_t, int, uint8_t*);
CABM_DECL(int, vaccinate, cabm_t*, int, const uint32_t*, size_t, uint8_t*); CABM_DECL(int, quarantine, cabm_t*, const uint32_t*, size_t, int, int);
CABM_DECL(int, ledger_table, cabm_t*, uint64_t*, size_t); CABM_DECL(int, cumulative_outcomes, cabm_t*, uint64_t*);
CABM_DECL(int, sync_histories_from, cabm_t*, size_t, cabm_infection*, size_t, size_t*);
CABM_DECL(int, detected_today, cabm_t*, int, uint32_t*, size_t, size_t*);
CABM_DECL(int, diagnostics, cabm_t*, int, uint64_t*); CABM_DECL(int, serosurvey, cabm_t*, int, const uint32_t*, size_t, uint64_t*);
CABM_DECL(int, snapshot, cabm_t*, cabm_snap_t**); CABM_DECL(int, restore, cabm_t*, const cabm_snap_t*); CABM_DECL(int, snap_free, cabm_snap_t*);
}

using namespace std;

extern gsl_rng* RNG;                                   // simulator.h:53-55 (defined by the including driver)
void cabm_facade_fill_params(const Parameters* par, cabm_params* out, std::vector<double>& pfd, std::vector<double>& imr);

const Parameters* Community::_par = nullptr;
const Parameters* Person::_par = nullptr;
size_t Person::NEXT_ID = 0;
size_t Location::NEXT_ID = 0;
bool Person::_stale = false;
Community* Person::_community = nullptr;

// ---------------------------------------------------------------------------------------------------------------------
// lifecycle
// ---------------------------------------------------------------------------------------------------------------------
// CABM_FACADE_TIMING=1: wall-clock seconds per facade entry point, one line on stderr when the community is destroyed
namespace {
struct FtSlot { const char* name; double s = 0; unsigned long n = 0; };
FtSlot ft_slots[] = { {"load_locations"}, {"load_population"}, {"upload"}, {"tick_device"}, {"tick_tracing"}, {"tick_vaccinate"}, {"infect"},
                      {"refresh_people"}, {"ledger_table"}, {"diagnostics"}, {"serosurvey_list"} };
enum { FT_LOADLOC, FT_LOADPOP, FT_UPLOAD, FT_TICK, FT_TRACE, FT_VAX, FT_INFECT, FT_REFRESH, FT_LEDGER, FT_DIAG, FT_SERO, FT_N };
const bool ft_on = getenv("CABM_FACADE_TIMING") != nullptr;
struct Ft {
    int k; std::chrono::steady_clock::time_point t0;
    explicit Ft(int k_) : k(k_) { if (ft_on) t0 = std::chrono::steady_clock::now(); }
    ~Ft() { if (ft_on) { ft_slots[k].s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); ft_slots[k].n++; } }
};
int ft_cache[2] = { -1, -1 };              // loadLocations / loadPopulation: 1 = rows from the binary cache, 0 = from the text, -1 = not called
void ft_report() {
    if (!ft_on) return;
    cerr << "facade_pop_cache {\"locations\": " << ft_cache[0] << ", \"population\": " << ft_cache[1] << "}" << endl;
    cerr << "facade_timing {";
    for (int k = 0; k < FT_N; ++k) cerr << (k ? ", " : "") << '"' << ft_slots[k].name << "\": [" << ft_slots[k].s << ", " << ft_slots[k].n << "]";
    cerr << "}" << endl;
}
}

Community::Community(const Parameters* parameters, Date* date)
    : _date(date), _day(0), cmty_ledger(nullptr), vac_campaign(nullptr), _h(nullptr), _uploaded(false), _ti_dirty(true), _epoch(1),
      _ledger_epoch(0), _diag_epoch(0), _diag_day(-1), _log_read(0) {
    _par = parameters;
    cmty_ledger = new CommunityLedger(parameters);
    cmty_ledger->_owner = this;
    _personAgeCohort.resize(NUM_AGE_CLASSES);
    for (int lt = 0; lt < (int) NUM_OF_LOCATION_TYPES; ++lt) _location_map[(LocationType) lt] = {};
    Person::_community = this;
    Person::_stale = false;
    const int rc = CABM_FN(create)(&_h);
    if (rc != 0) { cerr << "ERROR: cabm_create failed (rc " << rc << "): the CUDA library needs a visible B200" << endl; exit(-1); }
}

Community::~Community() {
    ft_report();
    if (_h) CABM_FN(destroy)(_h);
    if (_date) delete _date;
    if (cmty_ledger) delete cmty_ledger;
    if (vac_campaign) delete vac_campaign;
    Person::reset_ID_counter();                           // ids restart per community (Community.cpp:131)
    for (Person* p : _people) delete p;
    Location::reset_ID_counter();
    for (Location* l : _location) delete l;
    if (Person::_community == this) Person::_community = nullptr;
}

void Community::fail(const char* what, int rc) {
    cerr << "ERROR: " << what << " failed (rc " << rc << "): " << (_h ? CABM_FN(last_error)(_h) : "") << endl;
    exit(-1);                                             // the reference's convention (e.g. Community.cpp:80-81)
}

CommunityLedger::CommunityLedger(const CommunityLedger& o) : _owner(nullptr), _snap(o._snap), _timedInterventions(o._timedInterventions) {
    if (o._owner) _snap = o._owner->take_snapshot();      // copy of the live ledger = snapshot of the device state
}

std::shared_ptr<cabm_snap> Community::take_snapshot() {
    flush_pending();
    cabm_snap_t* s = nullptr;
    const int rc = CABM_FN(snapshot)(handle(), &s);
    if (rc != 0) fail("cabm_snapshot", rc);
    return std::shared_ptr<cabm_snap>(s, [](cabm_snap* p) { if (p) CABM_FN(snap_free)(p); });
}

void Community::load_from_cache(CommunityLedger* cache_ledger, Date* cache_date, map<int, vector<Person*>>, Vac_Campaign* cache_vc) {
    if (not cache_ledger->_snap) { cerr << "ERROR: load_from_cache needs a ledger copied from a live community" << endl; exit(-1); }
    const int rc = CABM_FN(restore)(handle(), cache_ledger->_snap.get());
    if (rc != 0) fail("cabm_restore", rc);
    cmty_ledger->_timedInterventions = cache_ledger->_timedInterventions; _ti_dirty = true;
    if (vac_campaign) delete vac_campaign;
    vac_campaign = cache_vc ? cache_vc->quick_cache() : nullptr;
    if (_date) delete _date;
    _date = new Date(*cache_date);
    // host mirrors are rebuilt from the restored device log; vaccination histories made after the snapshot are rolled back
    const int day = _date->day();
    for (Person* p : _people) {
        for (Infection* i : p->infectionHistory) delete i;
        p->infectionHistory.clear();
        while (not p->vaccineHistory.empty() and p->vaccineHistory.back() >= day) p->vaccineHistory.pop_back();
        while (not p->quarantineHistory.empty() and p->quarantineHistory.back().first >= day) p->quarantineHistory.pop_back();
    }
    _log_read = 0; _pending_doses.clear(); _pending_quar.clear();
    for (auto& q : _pending_inf) delete q.placeholder;      // (their owners' histories are cleared just above)
    _pending_inf.clear();
    device_changed();
}

void Community::device_changed() { ++_epoch; Person::_stale = true; }

// ---------------------------------------------------------------------------------------------------------------------
// upload on first use: parameters (drivers still edit Parameters after build_community, exp/active-vac/main.cpp:598-605)
// and the population as structure-of-arrays + CSR
// ---------------------------------------------------------------------------------------------------------------------
cabm* Community::handle() {
    if (_uploaded) return _h;
    Ft ft(FT_UPLOAD);
    cabm_params par; vector<double> pfd, imr;
    cabm_facade_fill_params(_par, &par, pfd, imr);
    int rc = CABM_FN(set_params)(_h, &par);
    if (rc != 0) fail("cabm_set_params", rc);

    const size_t N = _people.size(), L = _location.size();
    vector<uint8_t> loc_type(L), loc_ess(L), loc_pub(L); vector<float> loc_risk(L); vector<int32_t> loc_hosp(L);
    vector<int64_t> nbr_off(L + 1, 0), roster_off(L + 1, 0); vector<int32_t> nbr, roster;
    for (size_t l = 0; l < L; ++l) {
        Location* loc = _location[l];
        loc_type[l] = (uint8_t) loc->getType(); loc_ess[l] = loc->isEssential() ? 1 : 0; loc_pub[l] = (uint8_t) loc->getPublicTransmissionRisk();
        loc_risk[l] = loc->getRiskiness(); loc_hosp[l] = loc->getHospital() ? loc->getHospital()->getID() : -1;
        for (Location* n : loc->neighbors()) nbr.push_back(n->getID());              // std::set order = ascending id
        nbr_off[l + 1] = (int64_t) nbr.size();
        for (Person* p : loc->people()) roster.push_back(p->getID());
        roster_off[l + 1] = (int64_t) roster.size();
    }
    size_t n_patron = 1;
    for (Person* p : _people) n_patron = max(n_patron, p->patronized_locs.size());
    vector<int32_t> home(N), day(N), patron(N * n_patron, -1), nat_dur(N); vector<uint8_t> age(N), sex(N), com(N), ltc(N);
    vector<double> nat_eff(N); vector<float> xprot(N);
    for (size_t i = 0; i < N; ++i) {
        Person* p = _people[i];
        home[i] = p->home_loc->getID(); day[i] = p->day_loc ? p->day_loc->getID() : -1;
        age[i] = (uint8_t) p->age; sex[i] = (uint8_t) p->sex; com[i] = (uint8_t) p->comorbidity; ltc[i] = p->long_term_care ? 1 : 0;
        for (size_t k = 0; k < p->patronized_locs.size(); ++k) patron[i * n_patron + k] = p->patronized_locs[k]->getID();
        nat_eff[i] = p->startingNaturalEfficacy; xprot[i] = p->crossProtectionProbability; nat_dur[i] = p->naturalImmunityDuration;
    }
    cabm_population pop; memset(&pop, 0, sizeof(pop));
    pop.n_people = (uint32_t) N; pop.n_loc = (uint32_t) L;
    pop.loc_type = loc_type.data(); pop.loc_essential = loc_ess.data(); pop.loc_pubrisk = loc_pub.data(); pop.loc_risk = loc_risk.data();
    pop.loc_hospital = loc_hosp.data(); pop.nbr_off = nbr_off.data(); pop.nbr = nbr.data(); pop.roster_off = roster_off.data(); pop.roster = roster.data();
    pop.home = home.data(); pop.day = day.data(); pop.age = age.data(); pop.sex = sex.data(); pop.comorbid = com.data(); pop.ltc = ltc.data();
    pop.patron = patron.data(); pop.n_patron = (uint32_t) n_patron;
    pop.starting_natural_efficacy = nat_eff.data(); pop.cross_protection_probability = xprot.data(); pop.natural_immunity_duration = nat_dur.data();
    rc = CABM_FN(load_population)(_h, &pop);
    if (rc != 0) fail("cabm_load_population", rc);
    _uploaded = true;
    push_interventions();
    return _h;
}

void Community::push_interventions() {
    if (not _uploaded or not _ti_dirty) return;
    const TimedIntervention order[3] = { SCHOOL_CLOSURE, NONESSENTIAL_BUSINESS_CLOSURE, SOCIAL_DISTANCING };   // CABM_TI_* order
    for (int k = 0; k < 3; ++k) {
        vector<double> v = cmty_ledger->_timedInterventions[order[k]];
        v.resize(_par->runLength, v.empty() ? 0.0 : v.back());
        const int rc = CABM_FN(set_timed_intervention)(_h, k, v.data(), v.size());
        if (rc != 0) fail("cabm_set_timed_intervention", rc);
    }
    _ti_dirty = false;
}

// ---------------------------------------------------------------------------------------------------------------------
// loaders: the reference's five text formats (Community.cpp:174-524)
// ---------------------------------------------------------------------------------------------------------------------
static double pixel_of(double coord) { return floor(coord / 0.01) * 0.01 + 0.005; }           // Community.cpp:396-398

bool Community::loadLocations(string locationFilename, string networkFilename) {
    Ft ft(FT_LOADLOC);
    // text (or its binary cache, pop_cache.h) -> rows, then the proxies in row order
    cabm_pop::LocRows R;
    const int type_codes[5] = { (int) HOUSE, (int) WORK, (int) SCHOOL, (int) HOSPITAL, (int) NURSINGHOME };
    if (!cabm_pop::location_rows(locationFilename, networkFilename, type_codes, R, &_loc_from_cache)) return false;
    ft_cache[0] = _loc_from_cache ? 1 : 0;
    _location.clear();
    const size_t L = R.x.size();
    _location.reserve(L);
    for (size_t i = 0; i < L; ++i) {
        const LocationType lt = (LocationType) R.type[i];
        const unsigned int fl = R.flags[i];
        Location* loc = new Location();
        loc->setX(R.x[i]); loc->setY(R.y[i]); loc->setType(lt); loc->setEssential((fl & 2u) != 0);
        loc->setPixel(pixel_of(R.x[i]), pixel_of(R.y[i]));
        _pixelMap[{ loc->getXPixel(), loc->getYPixel() }].push_back(loc);
        if ((fl & 1u) and R.compliance[i] >= 0) loc->setRiskiness(1.0 - R.compliance[i]);
        else if (lt == HOUSE) loc->setRiskiness(gsl_rng_uniform(RNG));                     // Community.cpp:461-464
        const unsigned int pub = (fl >> 2) & 3u;
        const PublicTransmissionType pt = pub == 2 ? HIGH_PUBLIC_TRANSMISSION : pub == 1 ? LOW_PUBLIC_TRANSMISSION : NO_PUBLIC_TRANSMISSION;
        if (pt != NO_PUBLIC_TRANSMISSION) _public_locations.push_back(loc);
        loc->setPublicTransmissionRisk(pt);
        _location.push_back(loc);
        _location_map[lt].insert(loc);
    }
    for (size_t i = 0; i < L; ++i) {
        const int hf = R.hfid[i];
        if (hf < 0) continue;
        if (hf >= (int) L or _location[hf]->getType() != HOSPITAL) { cerr << "ERROR: hospital id " << hf << " is not a hospital" << endl; return false; }
        _location[i]->setHospital(_location[hf]);
    }
    for (size_t e = 0; e + 1 < R.net.size(); e += 2) {
        const int a = R.net[e], b = R.net[e + 1];
        assert(a >= 0 and b >= 0 and a < (int) L and b < (int) L);
        _location[a]->addNeighbor(_location[b]); _location[b]->addNeighbor(_location[a]);
    }
    return true;
}

bool Community::loadPopulation(string populationFilename, string comorbidityFilename, string publicActivityFilename, string immunityFilename) {
    Ft ft(FT_LOADPOP);
    cabm_pop::PopRows R;
    if (!cabm_pop::population_rows(populationFilename, comorbidityFilename, publicActivityFilename, R, &_pop_from_cache)) return false;
    ft_cache[1] = _pop_from_cache ? 1 : 0;
    const size_t N = R.hid.size();
    _people.reserve(_people.size() + N);
    for (size_t i = 0; i < N; ++i) {
        const int hid = R.hid[i], did = R.did[i];
        if ((int) i != (int) _people.size()) { cerr << "ERROR: Person ID's must be sequential integers starting at 0" << endl; return false; }
        assert(hid >= 0 and (int) _location.size() > hid and (int) _location.size() > did);
        Person* p = new Person();
        _people.push_back(p);
        p->setAge(R.age[i]); p->setSex((SexType) R.sex[i]);
        // load-time immunity attributes, drawn exactly as the reference does (Community.cpp:212-215) with its own Parameters methods
        p->setStartingNaturalEfficacy(_par->sampleStartingNaturalEfficacy(RNG));
        p->setImmunityQuantile(gsl_rng_uniform(RNG));
        p->setNaturalImmunityDuration(_par->immunityDuration(p->getImmunityQuantile(), p->getStartingNaturalEfficacy()));
        p->setCrossProtectionProbability(gsl_rng_uniform(RNG));
        p->setHomeLoc(_location[hid]); _location[hid]->addPerson(p);
        if (did >= 0) { p->setDayLoc(_location[did]); _location[did]->addPerson(p); }
        if (_location[hid]->getType() == NURSINGHOME) p->setLongTermCare(true);
    }
    for (const int pid : R.comorbid) getPersonByID(pid)->setComorbidity(COMORBID);
    for (size_t r = 0; r < R.pub_pid.size(); ++r) {
        Person* p = getPersonByID(R.pub_pid[r]);
        if (p->getAge() < 18 or p->getLongTermCare()) continue;                          // Community.cpp:286-288
        assert(R.pub_off[r] <= R.pub_off[r + 1] and R.pub_off[r + 1] <= (int64_t) R.pub_loc.size());
        for (int64_t k = R.pub_off[r]; k < R.pub_off[r + 1]; ++k) {
            const int locid = R.pub_loc[k];
            assert(locid >= 0 and (int) _location.size() > locid);
            Location* loc = _location[locid];
            assert(loc->getPublicTransmissionRisk() == LOW_PUBLIC_TRANSMISSION or loc->getPublicTransmissionRisk() == HIGH_PUBLIC_TRANSMISSION);
            p->addPatronizedLocation(loc);
        }
    }
    if (immunityFilename.length() > 0) { cerr << "ERROR: Reading in immunity file not currently supported." << endl; return false; }
    _personAgeCohort.clear(); _personAgeCohort.resize(NUM_AGE_CLASSES);
    for (Person* p : _people) { assert(p->getAge() >= 0 and p->getAge() < NUM_AGE_CLASSES); _personAgeCohort[p->getAge()].push_back(p); }
    _pediatric.clear();
    for (int a = 0; a <= 17; ++a) _pediatric.insert(_pediatric.end(), _personAgeCohort[a].begin(), _personAgeCohort[a].end());
    return true;
}

Person* Community::getPersonByID(int id) {
    if (id < 0 or id >= (int) _people.size()) { cerr << "ERROR: failed to find person with id " << id << " max: " << getNumPeople() << endl; exit(-1); }
    return _people[id];
}
Location* Community::getLocationByID(int id) {
    if (id < 0 or id >= (int) _location.size()) { cerr << "ERROR: failed to find location with id " << id << " max: " << getNumLocations() << endl; exit(-1); }
    return _location[id];
}

// ---------------------------------------------------------------------------------------------------------------------
// the day: Community::tick (Community.cpp:1147-1192)
// ---------------------------------------------------------------------------------------------------------------------
void Community::tick() {
    _day = _date->day();
    cabm* h = handle();
    flush_pending();
    push_interventions();
    { Ft ft(FT_TICK);
      const int rc = CABM_FN(tick)(h, _day, (int) _date->julianDay(), _date->isWeekday() ? 1 : 0);
      if (rc != 0) fail("cabm_tick", rc); }
    device_changed();

    // contact tracing to the configured depth from today's reported cases, then reactive scheduling and the day's doses
    vector<set<Person*, PerPtrComp>> tracedContactsByDepth;
    { Ft ft(FT_TRACE); tracedContactsByDepth = traceForwardContacts(); }
    if (vac_campaign) {
        Ft ft(FT_VAX);
        vac_campaign->reactive_strategy(_day, tracedContactsByDepth, this);
        vaccinate();
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// contact tracing: Community::traceForwardContacts (Community.cpp:1054-1145). The primary cases come from the device
// (cabm_detected_today: alive, latest infection reported today, ascending ids -- the order in which the reference walks
// _people and consumes RNG); the depth-limited search over household / neighbours / day location is the reference's
// sequential logic on the static proxies; quarantines are written through (Person::scheduleQuarantine -> cabm_quarantine).
// Difference: a hospital's roster here holds its staff only (admitted patients are device state), so the day-location
// contacts of hospital staff are drawn among staff.
// ---------------------------------------------------------------------------------------------------------------------
static void insert_new_living(set<Person*, PerPtrComp>& into, const vector<Person*>& from, set<Person*, PerPtrComp>& seen, int day) {
    for (Person* p : from) if (p->isAlive(day) and seen.insert(p).second) into.insert(p);
}

vector<set<Person*, PerPtrComp>> Community::traceForwardContacts() {
    vector<set<Person*, PerPtrComp>> traced(_par->contactTracingDepth + 1);
    if (_day < _par->beginContactTracing) return traced;
    size_t n = 0;
    int rc = CABM_FN(detected_today)(handle(), _day, nullptr, 0, &n);
    if (rc != 0) fail("cabm_detected_today", rc);
    vector<uint32_t> ids(n);
    if (n) { rc = CABM_FN(detected_today)(_h, _day, ids.data(), n, &n); if (rc != 0) fail("cabm_detected_today", rc); }
    set<Person*, PerPtrComp> cases;
    for (uint32_t id : ids) if (gsl_rng_uniform(RNG) < _par->contactTracingCoverage) cases.insert(_people[id]);
    traced[0] = cases;

    set<Person*, PerPtrComp> all_traced, all_interviewed;
    auto pick = [&](vector<Person*> pool, LocationType ev_of) {                  // Poisson(EV) contacts recalled out of `pool`
        const size_t k = gsl_ran_poisson(RNG, _par->contactTracingEV[ev_of]);
        return k <= pool.size() ? covid::util::choose_k(RNG, pool, k) : pool;
    };
    for (size_t depth = 0; depth < _par->contactTracingDepth; ++depth) {
        const set<Person*, PerPtrComp> interview = depth == 0 ? cases : traced[depth];
        for (Person* p : interview) {
            if (not (p->isAlive(_day) and all_interviewed.insert(p).second)) continue;
            Location* home = p->getHomeLoc();
            if (home->getType() == NURSINGHOME) {
                insert_new_living(traced[depth + 1], pick(home->getPeople(), NURSINGHOME), all_traced, _day);
            } else {
                insert_new_living(traced[depth + 1], home->getPeople(), all_traced, _day);
                vector<Person*> neighbours;
                for (Location* nb : home->neighbors()) for (Person* q : nb->people()) neighbours.push_back(q);
                insert_new_living(traced[depth + 1], pick(neighbours, HOUSE), all_traced, _day);
                vector<Person*> day_contacts;
                if (p->getDayLoc()) day_contacts = pick(p->getDayLoc()->getPeople(), p->getDayLoc()->getType());
                insert_new_living(traced[depth + 1], day_contacts, all_traced, _day);
            }
        }
    }
    for (size_t depth = 0; depth < traced.size(); ++depth)
        for (Person* p : traced[depth])
            if (not p->isQuarantining(_day) and gsl_rng_uniform(RNG) < _par->quarantineProbability[depth]) p->scheduleQuarantine(_day, _par->quarantineDuration);
    return traced;
}

Infection* Community::infect(int id, StrainType strain) {                        // Community.cpp:550-553
    return getPersonByID(id)->infect(this, _date, strain);
}

// Community::infect / Person::infect without a source (simulator.h:102,113,126,217) are QUEUED: seed_epidemic calls them once
// per imported case, hundreds of times a day, and a device round trip per case cost 25 times the day itself. The queue is
// resolved by one cabm_infect_batch (the same calls, in order, with their own sequence numbers) before anything reads or
// changes the device state. Until then the returned Infection* is provisional: a call that turns out not to take (the person
// was protected or already infected) has its placeholder removed again.
void Community::queue_infect(int pid, StrainType strain, int day, Infection* placeholder) {
    flush_doses();                                        // doses given before this call must be visible to it
    flush_quarantines();
    if (not _pending_inf.empty() and day != _pending_inf_day) flush_infects();
    _pending_inf_day = day;
    _pending_inf.push_back({ (unsigned int) pid, (unsigned char) strain, placeholder });
    device_changed();                                     // mirrors and caches are stale: the next reader resolves the queue
}

void Community::flush_infects() {
    if (_pending_inf.empty()) return;
    Ft ft(FT_INFECT);
    cabm* h = handle();
    const size_t n = _pending_inf.size();
    vector<uint32_t> pids(n); vector<uint8_t> strains(n), took(n, 0);
    for (size_t k = 0; k < n; ++k) { pids[k] = _pending_inf[k].pid; strains[k] = _pending_inf[k].strain; }
    const int rc = CABM_FN(infect_batch)(h, pids.data(), strains.data(), n, _pending_inf_day, took.data());
    if (rc != 0) fail("cabm_infect_batch", rc);
    for (size_t k = 0; k < n; ++k) {
        if (took[k]) continue;
        vector<Infection*>& hist = _people[pids[k]]->infectionHistory;
        for (size_t i = hist.size(); i-- > 0;) if (hist[i] == _pending_inf[k].placeholder) { hist.erase(hist.begin() + i); break; }
        delete _pending_inf[k].placeholder;
    }
    _pending_inf.clear();
    device_changed();
}

void Community::flush_pending() { flush_doses(); flush_infects(); flush_quarantines(); }

// Person::infect(community, source, date, sourceloc, strain) with no source is what seed_epidemic and Community::infect
// call (simulator.h:102,113,126,217); transmission between people never comes through here -- it happens inside cabm_tick.
Infection* Person::infect(Community* community, Person* source, const Date* date, Location*, StrainType strain, bool) {
    if (source) { cerr << "ERROR: person-to-person infection is resolved on the device, not through Person::infect" << endl; exit(-1); }
    if (strain == NUM_OF_STRAIN_TYPES) strain = WILDTYPE;
    Infection* placeholder = new Infection();             // completed from the device log at the next refresh, or withdrawn if the call does not take
    placeholder->infectedBegin = date->day(); placeholder->strain = strain; placeholder->infectionOwner = this;
    infectionHistory.push_back(placeholder);
    community->queue_infect(getID(), strain, date->day(), placeholder);
    return placeholder;
}

// pull the records appended to the device's infection log since the last read and rebuild the mirrors
void Person::refresh() { _stale = false; if (_community) _community->refresh_people(); }

void Community::refresh_people() {
    flush_pending();
    Ft ft(FT_REFRESH);
    Person::_stale = false;
    if (not _uploaded) return;
    size_t avail = 0;
    int rc = CABM_FN(sync_histories_from)(_h, _log_read, nullptr, 0, &avail);
    if (rc != 0) fail("cabm_sync_histories_from", rc);
    if (avail == 0) return;
    vector<cabm_infection> rec(avail);
    rc = CABM_FN(sync_histories_from)(_h, _log_read, rec.data(), avail, &avail);
    if (rc != 0) fail("cabm_sync_histories_from", rc);
    auto day_of = [](int16_t v) { return v == 32767 ? INT_MAX : v == -32768 ? INT_MIN : (int) v; };
    for (const cabm_infection& r : rec) {
        Person* p = _people[r.person];
        Infection* inf = nullptr;
        for (Infection* cand : p->infectionHistory) if (not cand->_filled) { inf = cand; break; }      // placeholder made by infect()
        if (!inf) { inf = new Infection(); p->infectionHistory.push_back(inf); }
        inf->_filled = true;
        inf->infectionOwner = p;
        inf->infectedBegin = day_of(r.t[0]); inf->infectiousBegin = day_of(r.t[1]); inf->infectiousEnd = day_of(r.t[2]);
        inf->symptomBegin = day_of(r.t[3]); inf->symptomEnd = day_of(r.t[4]); inf->severeBegin = day_of(r.t[5]); inf->severeEnd = day_of(r.t[6]);
        inf->hospitalizedBegin = day_of(r.t[7]); inf->criticalBegin = day_of(r.t[8]); inf->criticalEnd = day_of(r.t[9]);
        inf->icuBegin = day_of(r.t[10]); inf->deathTime = day_of(r.t[11]);
        inf->strain = (StrainType) r.strain; inf->relInfectiousness = r.rel_infectiousness;
        inf->_has_detection = r.detected_state < NUM_OF_OUTCOME_TYPES;
        if (inf->_has_detection) inf->_detection = Detection((OutcomeType) r.detected_state, day_of(r.report_day));
        inf->infectedPlace = r.infected_place >= 0 ? _location[r.infected_place] : nullptr;
        inf->infectedBy = nullptr;
        if (r.infected_by >= 0) {
            Person* src = _people[r.infected_by];
            inf->infectedBy = src;
            for (Infection* s : src->infectionHistory)                                                 // log_transmission (Community.cpp:819)
                if (s->_filled and s->isInfectious(inf->infectedBegin)) { s->infections_caused.push_back(inf); break; }
        }
    }
    _log_read += avail;
}

// ---------------------------------------------------------------------------------------------------------------------
// vaccination: the campaign is the reference's own Vac_Campaign object; this is the control flow of
// Community::vaccinate (Community.cpp:556-629) with the dose state written through to the device in one batch
// ---------------------------------------------------------------------------------------------------------------------
void Community::vaccinate() {
    while (vac_campaign->add_new_eligible_people(_day)) {}
    if (not vac_campaign->get_all_doses_available(_day)) return;
    vector<Eligibility_Group*> urg_revax = vac_campaign->init_new_eligible_groups(_day);
    vector<Eligibility_Group*> std_revax = vac_campaign->init_new_eligible_groups(_day);
    vector<map<int, map<int, int>>> sampled = vac_campaign->multinomially_distribute_pooled_doses(_day);
    const vector<int> bins = vac_campaign->get_unique_age_bins();
    for (int dose = 0; dose < _par->numVaccineDoses; ++dose) {
        for (int bin : bins) {
            for (Vaccinee* v = vac_campaign->next_vaccinee(_day, dose, bin, sampled); v; v = vac_campaign->next_vaccinee(_day, dose, bin, sampled)) {
                const Person* p = v->get_person();
                const bool eligible = p->getNumVaccinations() == dose and p->isSeroEligible() and p->isInfEligible(_day);
                if (eligible and vac_campaign->vaccinate(v, _day)) {
                    vac_campaign->tally_dose(_day, dose, bin, v, sampled);
                    if (dose + 1 < _par->numVaccineDoses) vac_campaign->assign_vaccinee_for_revaccination(v, dose, bin, std_revax, urg_revax);
                }
                vac_campaign->remove_from_pool(dose, bin, v);
                delete v;
            }
        }
    }
    for (int bin : bins) for (int dose = 0; dose < _par->numVaccineDoses; ++dose) vac_campaign->rollover_unused_doses(_day, dose, bin);
    vac_campaign->schedule_revaccinations(urg_revax, std_revax);
    flush_pending();
}

bool Person::vaccinate(int time) {                                                // Person.cpp:611-647 (leaky immunity only, :622)
    if (not isAlive(time)) return false;
    vaccineHistory.push_back(time);
    _community->queue_dose(getID());
    return true;
}
void Community::queue_dose(int pid) { flush_infects(); _pending_doses.push_back((unsigned int) pid); }

void Community::flush_doses() {
    if (_pending_doses.empty()) return;
    vector<uint8_t> ok(_pending_doses.size());
    const int rc = CABM_FN(vaccinate)(handle(), _day, _pending_doses.data(), _pending_doses.size(), ok.data());
    if (rc != 0) fail("cabm_vaccinate", rc);
    for (uint8_t o : ok) if (!o) { cerr << "ERROR: the device refused a dose the host mirror allowed (stale death state)" << endl; exit(-1); }
    _pending_doses.clear();
    device_changed();
}

bool Person::isImmuneState(ImmuneStateType is) const {
    const bool nat = hasBeenInfected(), vax = isVaccinated();
    const ImmuneStateType mine = nat ? (vax ? NATURAL_AND_VACCINATED : NATURAL) : (vax ? VACCINATED : NAIVE);
    return is == mine;
}

bool Person::isSeroEligible() const {                                             // Person.cpp:562-583
    const VaccineSeroConstraint vsc = _par->vaccineSeroConstraint;
    if (vsc == VACCINATE_ALL_SERO_STATUSES) return true;
    bool seropos = not isImmuneState(NAIVE);
    if ((seropos and _par->seroTestFalseNeg > gsl_rng_uniform(RNG)) or (not seropos and _par->seroTestFalsePos > gsl_rng_uniform(RNG))) seropos = not seropos;
    return vsc == VACCINATE_SEROPOSITIVE_ONLY ? seropos : not seropos;
}

bool Person::isInfEligible(int today) const {                                      // Person.cpp:586-608
    const VaccineInfConstraint vic = _par->vaccineInfConstraint;
    if (vic == VACCINATE_ALL_INF_STATUSES) return true;
    const bool prior_inf = hasBeenInfected();
    bool prior_case = false;
    for (Infection* inf : infectionHistory) if (inf->isDetected() and inf->getDetection()->reported_time <= today) { prior_case = true; break; }
    switch (vic) {
        case VACCINATE_NAIVE_ONLY:    return not prior_inf;
        case VACCINATE_INF_ONLY:      return prior_inf;
        case VACCINATE_NON_CASE_ONLY: return not prior_case;
        case VACCINATE_CASE_ONLY:     return prior_case;
        default:                      return true;
    }
}

void Person::scheduleQuarantine(int start, int quarantineDuration) {               // Person.cpp:649-651
    quarantineHistory.emplace_back(start, start + quarantineDuration);
    _community->push_quarantine(getID(), start, start + quarantineDuration);
}
// contact tracing quarantines hundreds to thousands of people a day, all for the same interval (Community.cpp:1135): queued and
// written to the device in one call per interval
void Community::push_quarantine(int pid, int from, int until) {
    if (not _pending_quar.empty() and (from != _pq_from or until != _pq_until)) flush_quarantines();
    _pq_from = from; _pq_until = until;
    _pending_quar.push_back((unsigned int) pid);             // no epoch bump here: a quarantine changes no infection history, and the readers
}                                                            // that depend on it (tick, diagnostics, snapshot) flush first
void Community::flush_quarantines() {
    if (_pending_quar.empty()) return;
    const int rc = CABM_FN(quarantine)(handle(), _pending_quar.data(), _pending_quar.size(), _pq_from, _pq_until);
    if (rc != 0) fail("cabm_quarantine", rc);
    _pending_quar.clear();
    device_changed();
}

// ---------------------------------------------------------------------------------------------------------------------
// getters: ledger series (one device read per state change), cumulative outcomes, daily diagnostics
// ---------------------------------------------------------------------------------------------------------------------
std::vector<size_t> Community::series(int counter) {
    const size_t R = _par->runLength;
    flush_pending();
    if (not _uploaded) return std::vector<size_t>(R, 0);
    if (_ledger_epoch != _epoch) {
        Ft ft(FT_LEDGER);
        _ledger_cache.assign(R * CABM_N_COUNTERS, 0);
        const int rc = CABM_FN(ledger_table)(_h, (uint64_t*) _ledger_cache.data(), _ledger_cache.size());
        if (rc != 0) fail("cabm_ledger_table", rc);
        _ledger_epoch = _epoch;
    }
    return std::vector<size_t>(_ledger_cache.begin() + counter * R, _ledger_cache.begin() + (counter + 1) * R);
}

std::vector<size_t> Community::getNumNewInfectionsByLoc(std::string) { return std::vector<size_t>(_par->runLength, 0); }

std::vector<size_t> Community::getCumulIncidenceByOutcome() {
    uint64_t out[CABM_N_OUTCOMES] = { 0 };
    flush_pending();
    if (_uploaded) { const int rc = CABM_FN(cumulative_outcomes)(_h, out); if (rc != 0) fail("cabm_cumulative_outcomes", rc); }
    return std::vector<size_t>(out, out + CABM_N_OUTCOMES);
}

const std::vector<unsigned long long>& Community::diagnostics(int day) {
    flush_pending();
    if (_diag_epoch != _epoch or _diag_day != day) {
        Ft ft(FT_DIAG);
        _diag_cache.assign(CABM_N_DIAG, 0);
        const int rc = CABM_FN(diagnostics)(handle(), day, (uint64_t*) _diag_cache.data());
        if (rc != 0) fail("cabm_diagnostics", rc);
        _diag_epoch = _epoch; _diag_day = day;
    }
    return _diag_cache;
}

size_t Community::getNumInfected(int day)    { return _uploaded or _people.size() ? diagnostics(day)[CABM_D_INFECTED] : 0; }
size_t Community::getNumInfectious(int day)  { return diagnostics(day)[CABM_D_INFECTIOUS]; }
size_t Community::getNumSymptomatic(int day) { return diagnostics(day)[CABM_D_SYMPTOMATIC]; }
size_t Community::getNumNaive()              { return diagnostics(_day)[CABM_D_NAIVE]; }
double Community::getNumPeopleQuarantining(int day) { return (double) diagnostics(day)[CABM_D_QUARANTINING] / _people.size(); }

// Community::calculate_vax_stats (Community.cpp:1329-1423): the same ratios on the device counts
std::map<std::string, double> Community::calculate_vax_stats(int day) {
    const std::vector<unsigned long long>& c = diagnostics(day);
    const double n = (double) getNumPeople(), full = (double) c[CABM_D_FULLY_VAXD], unv = n - full;
    auto ve = [&](int bt, int uv) { return 1.0 - ((double) c[bt] / full) / ((double) c[uv] / unv); };
    std::map<std::string, double> m;
    m["coverage"] = full / n; m["dose_1"] = c[CABM_D_DOSE1] / n; m["dose_2"] = c[CABM_D_DOSE2] / n; m["dose_3"] = c[CABM_D_DOSE3] / n;
    m["VES"] = ve(CABM_D_BT_INF, CABM_D_UV_INF); m["VEP"] = ve(CABM_D_BT_DIS, CABM_D_UV_DIS);
    m["VEH"] = ve(CABM_D_BT_HOSP, CABM_D_UV_HOSP); m["VED"] = ve(CABM_D_BT_DEATH, CABM_D_UV_DEATH);
    m["breakthruRatio"] = (double) c[CABM_D_BT_REPORTED] / ((double) c[CABM_D_BT_REPORTED] + (double) c[CABM_D_UV_REPORTED]);
    m["vaxInfs"] = (double) c[CABM_D_BT_INF]; m["unvaxInfs"] = (double) c[CABM_D_UV_INF];
    m["vaxHosp"] = (double) c[CABM_D_BT_HOSP]; m["unvaxHosp"] = (double) c[CABM_D_UV_HOSP];
    return m;
}

// Community::doSerosurvey (Community.cpp:1224-1274). Everybody and the 0-17 cohort (the two calls of simulator.h:403-410)
// come out of the daily reduction; any other list goes to cabm_serosurvey.
double Community::doSerosurvey(const ImmuneStateType ist, const std::vector<Person*>& pop, int time) {
    if (ist != NATURAL) { cerr << "ERROR: Community::doSerosurvey() not yet implemented for this immunity type" << endl; exit(1); }
    if (pop.size() == _people.size()) return (double) diagnostics(time)[CABM_D_SEROPOS_ALL] / pop.size();
    if (pop.size() == _pediatric.size() and not pop.empty() and std::memcmp(pop.data(), _pediatric.data(), pop.size() * sizeof(Person*)) == 0)
        return (double) diagnostics(time)[CABM_D_SEROPOS_PED] / pop.size();
    Ft ft(FT_SERO);
    std::vector<uint32_t> ids(pop.size());
    for (size_t i = 0; i < pop.size(); ++i) ids[i] = (uint32_t) pop[i]->getID();
    uint64_t pos = 0;
    flush_pending();
    const int rc = CABM_FN(serosurvey)(handle(), time, ids.data(), ids.size(), &pos);
    if (rc != 0) fail("cabm_serosurvey", rc);
    return (double) pos / pop.size();
}

// ---------------------------------------------------------------------------------------------------------------------
// timed interventions (Community.h:396-425); the device copy is refreshed before the next tick
// ---------------------------------------------------------------------------------------------------------------------
double Community::social_distancing(int day) { return cmty_ledger->_timedInterventions[SOCIAL_DISTANCING][day]; }

void Community::updateTimedIntervention(TimedIntervention ti, size_t date, double val) {
    std::vector<double>& v = cmty_ledger->_timedInterventions.at(ti);
    const size_t n = v.size();
    v.resize(date); v.resize(n, val);
    _ti_dirty = true;
}
void Community::setSocialDistancingTimedIntervention(const std::vector<double>& vec) {
    std::vector<double>& v = cmty_ledger->_timedInterventions[SOCIAL_DISTANCING];
    v = vec; v.resize(_par->runLength, vec.back());
    _ti_dirty = true;
}
void Community::setSocialDistancingTimedIntervention(const std::vector<TimeSeriesAnchorPoint>& vec) {
    setSocialDistancingTimedIntervention(Date::linInterpolateTimeSeries(vec, _par->startJulianYear, _par->startDayOfYear));
}
void Community::_extendSocialDistancingTimedIntervention(double val) {
    std::vector<double>& v = cmty_ledger->_timedInterventions[SOCIAL_DISTANCING];
    v.resize(v.size() + _par->tuning_window, val);
    _ti_dirty = true;
}
void Community::_reviseSocialDistancingTimedIntervention(double val) {
    std::vector<double>& v = cmty_ledger->_timedInterventions[SOCIAL_DISTANCING];
    assert(v.size() >= _par->tuning_window);
    v.resize(v.size() - _par->tuning_window);
    _extendSocialDistancingTimedIntervention(val);
}

// ---------------------------------------------------------------------------------------------------------------------
// post-run analytics: every experiment driver asks for Rt by day of infection (exp/active-vac/main.cpp:482,749). They read
// the infector links of the device's infection log through the Person / Infection proxies.
// ---------------------------------------------------------------------------------------------------------------------
// Community.cpp:631-654: entry d = (infections that began on day d, mean number of infections each of them went on to cause)
vector<pair<size_t, double>> Community::getMeanNumSecondaryInfections() {
    vector<size_t> n(_par->runLength, 0), caused(_par->runLength, 0);
    for (Person* p : getPeople()) for (Infection* inf : p->getInfectionHistory()) {
        const int onset = inf->getInfectedTime();
        if (onset < 0 or onset >= (int) _par->runLength) continue;                        // historical infection
        n[onset]++; caused[onset] += inf->secondary_infection_tally();
    }
    vector<pair<size_t, double>> Rt(_par->runLength, { 0, 0.0 });
    for (size_t d = 0; d < Rt.size(); ++d) if (n[d]) Rt[d] = { n[d], (double) caused[d] / (double) n[d] };
    return Rt;
}

// Community.cpp:1425-1441: histogram of the number of secondary infections per infection, at least ten bins
vector<size_t> Community::generateOffspringDistribution() {
    vector<size_t> hist(10, 0);
    for (Person* p : getPeople()) for (Infection* inf : p->getInfectionHistory()) {
        const size_t k = inf->secondary_infection_tally();
        if (k >= hist.size()) hist.resize(k + 1, 0);
        hist[k]++;
    }
    return hist;
}

// Community.cpp:1281-1327: household secondary attack rate of the month two months back, called on the first of a month.
// An index case is somebody whose latest infection began between the start of the survey month and one survey-month length
// past today, and whose LAST-listed housemate was not infected in the ten days up to that infection (the reference's loop
// keeps the verdict of the last housemate it looks at; kept). Each index with at least one household member among the
// infections it caused contributes (such members) / (household size); the mean of those is returned.
double Community::getHouseholdSecondaryAttackRate(vector<Person*>& pop) {
    if (pop.empty()) pop = getPeople();
    if (not (_date->month() >= 3)) return 0.0;
    const int today = _date->day();
    const size_t jm = _date->julianMonth();
    const size_t survey_m = (int) jm - 2 >= 1 ? jm - 2 : jm + 10, between_m = (int) jm - 1 >= 1 ? jm - 1 : jm + 11;
    const vector<size_t>& dim = _date->isLeap() ? LEAP_DAYS_IN_MONTH : COMMON_DAYS_IN_MONTH;
    const int first = today - (int) dim.at(between_m - 1) - (int) dim.at(survey_m - 1), last = today + (int) dim.at(survey_m - 1);
    assert(first >= 0);
    vector<double> sar;
    for (Person* p : pop) {
        if (not p->hasBeenInfected()) continue;
        const int t = p->getInfectedTime();
        if (t < first or t > last) continue;
        bool index_case = false;
        for (Person* fam : p->getHomeLoc()->getPeople()) {
            if (fam == p) continue;
            index_case = not (fam->hasBeenInfected() and fam->getInfectedTime() >= t - 10 and fam->getInfectedTime() <= t);
        }
        if (not index_case) continue;
        int at_home = 0;
        for (Infection* sec : p->getInfection()->get_infections_caused()) if (sec->getInfectionOwner()->getHomeLoc() == p->getHomeLoc()) ++at_home;
        if (at_home > 0) sar.push_back((double) at_home / p->getHomeLoc()->getNumPeople());
    }
    return covid::util::mean(sar);
}
