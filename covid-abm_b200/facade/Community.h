// covid-abm_b200/facade/Community.h -- drop-in for the reference's Community.h (Community.h:23-493).
//
// Same public surface as the reference's `Community` for everything simulator.h, behavior_tuner.h and
// Vac_Campaign.cpp call, with the bodies forwarding to the C ABI of include/cabm.h. The agent and location state of
// the daily transmission cycle is owned by the GPU; this class owns the handle, the host proxies (Person, Location),
// the timed interventions and the vaccination campaign object (the reference's own, unmodified Vac_Campaign).
#ifndef CABM_FACADE_COMMUNITY_H
#define CABM_FACADE_COMMUNITY_H

#include <map>
#include <memory>
#include <set>
#include <string>
#include <utility>
#include <vector>
#include <cassert>
#include "Date.h"
#include "Location.h"
#include "Vac_Campaign.h"

class Person;
struct cabm;                    // opaque handle of include/cabm.h
struct cabm_snap;

// What `new CommunityLedger(*community->get_ledger())` has to capture (behavior_tuner.h:54): the reference copies its
// tally vectors and the hot-location calendar; here that state is on the device, so a copy of the live ledger takes
// a device snapshot (cabm_snapshot) and Community::load_from_cache restores it (cabm_restore).
class CommunityLedger {
    friend class Community;
  public:
    CommunityLedger(const Parameters* par) : _owner(nullptr), _timedInterventions(par->timedInterventions) {}
    CommunityLedger(const CommunityLedger& o);
    ~CommunityLedger() {}
  private:
    Community* _owner;                                   // set on the live ledger only
    std::shared_ptr<cabm_snap> _snap;                    // set on cached copies only
    std::map<TimedIntervention, std::vector<double>> _timedInterventions;
};

class Community {
  public:
    Community(const Parameters* parameters, Date* date);
    Community(const Community&) = delete;
    ~Community();

    Date* get_date() { return _date; }
    bool loadPopulation(std::string populationFilename, std::string comorbidityFilename = "", std::string publicActivityFilename = "",
                        std::string immunityFilename = "");
    bool loadLocations(std::string locationFilename, std::string networkFilename = "");
    size_t getNumPeople() const { return _people.size(); }
    size_t getNumLocations() const { return _location.size(); }
    const std::vector<Person*>& getPeople() const { return _people; }     // the reference returns a copy (Community.h:330); a reference binds wherever a copy did
    Person* getPersonByID(int id);
    Location* getLocationByID(int id);
    int getDay() { return _day; }

    Infection* infect(int id, StrainType strain);
    void tick();                                          // Community.cpp:1147-1192 -> cabm_tick + host-side campaign
    void vaccinate();                                     // Community.cpp:556-629, on the reference's own Vac_Campaign
    std::vector<std::set<Person*, PerPtrComp>> traceForwardContacts();   // Community.cpp:1054-1145, primary cases from the device
    double social_distancing(int day);

    // daily diagnostics of the caller, served by one device reduction per (day, state) (cabm_diagnostics)
    size_t getNumInfected(int day);
    size_t getNumInfectious(int day);
    size_t getNumSymptomatic(int day);
    size_t getNumNaive();
    double getNumPeopleQuarantining(int day);
    std::map<std::string, double> calculate_vax_stats(int day);
    // post-run analytics over the infection records (Community.cpp:631-654,1281-1327,1425-1441), on the host mirror of the device log
    std::vector<std::pair<size_t, double>> getMeanNumSecondaryInfections();
    std::vector<size_t> generateOffspringDistribution();
    double getHouseholdSecondaryAttackRate(std::vector<Person*>& pop);
    double doSerosurvey(const ImmuneStateType ist, const std::vector<Person*>& pop, int time);

    void setVac_Campaign(Vac_Campaign* vc) { vac_campaign = vc; }
    Vac_Campaign* getVac_Campaign() const { return vac_campaign; }

    // ledger getters (Community.h:353-375): by value, one device read of the whole table per (day, state)
    std::vector<size_t> getNumNewlyInfected()            { return series(0); }
    std::vector<size_t> getNumNewInfections(StrainType strain) { return series(16 + (int) strain); }
    std::vector<size_t> getNumNewInfectionsByLoc(std::string);                 // not tallied by the device path: zeros
    std::vector<size_t> getNumNewlySymptomatic()         { return series(1); }
    std::vector<size_t> getNumNewlySevere()              { return series(2); }
    std::vector<size_t> getNumNewlyCritical()            { return series(3); }
    std::vector<size_t> getNumNewlyDead()                { return series(4); }
    std::vector<size_t> getNumVaccinatedCases()          { return series(5); }
    std::vector<size_t> getNumSeverePrev()               { return series(6); }
    std::vector<size_t> getNumHospInc()                  { return series(7); }
    std::vector<size_t> getNumHospPrev()                 { return series(8); }
    std::vector<size_t> getNumIcuInc()                   { return series(9); }
    std::vector<size_t> getNumIcuPrev()                  { return series(10); }
    std::vector<size_t> getNumDetectedCasesOnset()       { return series(11); }
    std::vector<size_t> getNumDetectedCasesReport()      { return series(12); }
    std::vector<size_t> getNumDetectedHospitalizations() { return series(13); }
    std::vector<size_t> getNumDetectedDeathsOnset()      { return series(14); }
    std::vector<size_t> getNumDetectedDeathsReport()     { return series(15); }
    std::vector<size_t> getCumulIncidenceByOutcome();
    size_t getCumulIncidenceByOutcome(OutcomeType ot) { return getCumulIncidenceByOutcome()[ot]; }

    const std::vector<Location*> getLocations() const { return _location; }
    std::set<Location*, Location::LocPtrComp> getLocationsByType(LocationType lt) const { return _location_map.at(lt); }
    const std::vector<Person*>& getAgeCohort(unsigned int age) const { assert(age < _personAgeCohort.size()); return _personAgeCohort[age]; }
    std::vector<Location*> locsAtPixel(std::pair<double, double> px) { return _pixelMap[{ px.first, px.second }]; }
    std::vector<Location*> locsAtPixel(double xP, double yP) { return _pixelMap[{ xP, yP }]; }

    // timed interventions (Community.h:396-425); the device copy is refreshed before the next tick
    std::vector<double> getTimedIntervention(TimedIntervention ti) const { return cmty_ledger->_timedInterventions.at(ti); }
    double getTimedIntervention(TimedIntervention ti, size_t day) const { return cmty_ledger->_timedInterventions.at(ti)[day]; }
    void updateTimedIntervention(TimedIntervention ti, size_t date, double val);
    void setSocialDistancingTimedIntervention(const std::vector<double>& vec);
    void setSocialDistancingTimedIntervention(const std::vector<TimeSeriesAnchorPoint>& vec);
    void _clearSocialDistancingTimedIntervention() { cmty_ledger->_timedInterventions[SOCIAL_DISTANCING].clear(); _ti_dirty = true; }
    void _extendSocialDistancingTimedIntervention(double val);
    void _reviseSocialDistancingTimedIntervention(double val);

    CommunityLedger* get_ledger() const { return cmty_ledger; }
    void load_from_cache(CommunityLedger* cache_ledger, Date* cache_date, std::map<int, std::vector<Person*>> cache_hosp_ppl, Vac_Campaign* cache_vc);

    // ---- facade internals used by Person / CommunityLedger ----
    cabm* handle();                                       // uploads population and parameters on first use
    void device_changed();                                // bumps the state epoch: mirrors and caches are stale
    void queue_infect(int pid, StrainType strain, int day, Infection* placeholder);   // no-source infections are resolved in one batch (cabm_infect_batch)
    void queue_dose(int pid);
    void push_quarantine(int pid, int from, int until);
    void refresh_people();
    std::shared_ptr<cabm_snap> take_snapshot();

  protected:
    static const Parameters* _par;
    Date* _date;
    std::vector<Person*> _people;
    std::vector<std::vector<Person*>> _personAgeCohort;
    std::vector<Person*> _pediatric;                                            // cohorts 0..17 in simulate_epidemic's order (simulator.h:405-409)
    std::vector<Location*> _location;
    std::vector<Location*> _public_locations;
    std::map<LocationType, std::set<Location*, Location::LocPtrComp>> _location_map;
    std::map<std::pair<double, double>, std::vector<Location*>> _pixelMap;
    int _day;
    CommunityLedger* cmty_ledger;
    Vac_Campaign* vac_campaign;

  private:
    std::vector<size_t> series(int counter);
    const std::vector<unsigned long long>& diagnostics(int day);
    void flush_doses();
    void flush_infects();
    void flush_quarantines();
    void flush_pending();                                 // everything queued on the host reaches the device: doses, no-source infections, quarantines
    void push_interventions();
    void fail(const char* what, int rc);

    cabm* _h;
    bool _uploaded, _ti_dirty;
    bool _loc_from_cache = false, _pop_from_cache = false;   // the loaders read their rows from the binary cache (pop_cache.h, CABM_POP_CACHE)
    unsigned long long _epoch;                            // bumped whenever the device state changes
    unsigned long long _ledger_epoch, _diag_epoch; int _diag_day;
    std::vector<unsigned long long> _ledger_cache, _diag_cache;
    size_t _log_read;                                     // records of the device infection log already mirrored
    std::vector<unsigned int> _pending_doses;
    struct PendingInfect { unsigned int pid; unsigned char strain; Infection* placeholder; };
    std::vector<PendingInfect> _pending_inf; int _pending_inf_day = 0;
    std::vector<unsigned int> _pending_quar; int _pq_from = 0, _pq_until = 0;
};
#endif
