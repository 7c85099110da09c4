// covid-abm_b200/facade/Person.h -- drop-in for the reference's Person.h (Person.h:18-385).
//
// Host-side proxies. A Person keeps the static attributes the reference's callers read (id, age, sex, homes,
// comorbidity ...) and a *mirror* of the two histories that live on the GPU: infectionHistory is rebuilt from the
// device's append-only infection log (cabm_sync_histories_from) the first time anybody looks at it after the device
// state changed; vaccineHistory / quarantineHistory are written through (Person::vaccinate -> batched
// cabm_vaccinate, scheduleQuarantine -> cabm_quarantine). No disease logic runs here.
#ifndef CABM_FACADE_PERSON_H
#define CABM_FACADE_PERSON_H

#include <climits>
#include <vector>
#include "Parameters.h"
#include "Location.h"

class Community;
class Vaccinee;
class Person;

struct Detection {
    Detection() : detected_state(NUM_OF_OUTCOME_TYPES), reported_time(INT_MAX) {}
    Detection(OutcomeType ds, int rt) : detected_state(ds), reported_time(rt) {}
    OutcomeType detected_state;
    int reported_time;
};

// One entry of Person::infectionHistory; every field is final when the infection is created (the whole course is
// scheduled by Person::infect, Person.cpp:320-494), except infections_caused which grows as the log is read.
class Infection {
    friend class Person;
    friend class Community;
  public:
    Infection() : infectedBegin(INT_MAX), infectedPlace(nullptr), infectedBy(nullptr), infectionOwner(nullptr), infectiousBegin(INT_MAX),
                  infectiousEnd(INT_MIN), symptomBegin(INT_MAX), symptomEnd(INT_MIN), severeBegin(INT_MAX), severeEnd(INT_MIN),
                  hospitalizedBegin(INT_MAX), criticalBegin(INT_MAX), criticalEnd(INT_MIN), icuBegin(INT_MAX), deathTime(INT_MAX),
                  strain(WILDTYPE), relInfectiousness(1.0), _has_detection(false), _filled(false) {}

    bool isLocallyAcquired()      const { return infectedBy != nullptr; }
    Location* getInfectedPlace()  const { return infectedPlace; }
    void setInfectedPlace(Location* loc) { infectedPlace = loc; }
    Person* getInfectedBy()       const { return infectedBy; }
    void setInfectedBy(Person* p)       { infectedBy = p; }
    Person* getInfectionOwner()   const { return infectionOwner; }
    void setInfectionOwner(Person* p)   { infectionOwner = p; }

    int getInfectedTime()         const { return infectedBegin; }
    int getInfectiousTime()       const { return infectiousBegin; }
    int getInfectiousEndTime()    const { return infectiousEnd; }
    int getSymptomTime()          const { return symptomBegin; }
    int getSymptomEndTime()       const { return symptomEnd; }
    int getSevereTime()           const { return severeBegin; }
    int getSevereEndTime()        const { return severeEnd; }
    int getHospitalizedTime()     const { return hospitalizedBegin; }
    int getCriticalTime()         const { return criticalBegin; }
    int getCriticalEndTime()      const { return criticalEnd; }
    int getIcuTime()              const { return icuBegin; }
    int getDeathTime()            const { return deathTime; }
    StrainType getStrain()        const { return strain; }
    double getRelInfectiousness() const { return relInfectiousness; }

    bool infected()    const { return infectedBegin     != INT_MAX; }
    bool infectious()  const { return infectiousBegin   != INT_MAX; }
    bool symptomatic() const { return symptomBegin      != INT_MAX; }
    bool severe()      const { return severeBegin       != INT_MAX; }
    bool critical()    const { return criticalBegin     != INT_MAX; }
    bool hospital()    const { return hospitalizedBegin != INT_MAX; }
    bool icu()         const { return icuBegin          != INT_MAX; }
    bool fatal()       const { return deathTime         != INT_MAX; }

    bool isInfected(int now)    const { return infectedBegin <= now and now < infectiousEnd; }
    bool isInfectious(int now)  const { return infectiousBegin <= now and now < infectiousEnd; }
    bool isSymptomatic(int now) const { return symptomBegin <= now and now < symptomEnd; }
    bool isSevere(int now)      const { return severeBegin <= now and now < severeEnd; }
    bool isCritical(int now)    const { return criticalBegin <= now and now < criticalEnd; }
    bool inHospital(int now)    const { return hospitalizedBegin <= now and now < severeEnd; }
    bool inIcu(int now)         const { return icuBegin <= now and now < criticalEnd; }
    bool isDead(int now)        const { return deathTime <= now; }

    Detection* getDetection()        { return _has_detection ? &_detection : nullptr; }
    bool isDetected()          const { return _has_detection; }
    bool isDetectedOn(int now) const { return _has_detection and _detection.reported_time == now; }

    size_t secondary_infection_tally() const { return infections_caused.size(); }
    std::vector<Infection*> get_infections_caused() { return infections_caused; }
    std::vector<int> generation_times() const {
        std::vector<int> times;
        for (Infection* other : infections_caused) times.push_back(other->getInfectedTime() - getInfectedTime());
        return times;
    }

  private:
    int infectedBegin;
    Location* infectedPlace;
    Person* infectedBy;
    Person* infectionOwner;
    int infectiousBegin, infectiousEnd, symptomBegin, symptomEnd, severeBegin, severeEnd, hospitalizedBegin, criticalBegin, criticalEnd, icuBegin, deathTime;
    std::vector<Infection*> infections_caused;
    StrainType strain;
    double relInfectiousness;
    Detection _detection;
    bool _has_detection;
    bool _filled;                       // false while this is only the placeholder returned by infect() before the next log read
};

class Person {
    friend class Vaccinee;
    friend class Community;
  public:
    Person() : id(NEXT_ID++), home_loc(nullptr), day_loc(nullptr), age(-1), sex(UNKNOWN), long_term_care(false), comorbidity(HEALTHY),
               naturalImmunityDuration(0), startingNaturalEfficacy(0.0), immunityQuantile(0.0), crossProtectionProbability(0.0f) {}
    ~Person() { for (Infection* i : infectionHistory) delete i; }

    int getID() const { return (int) id; }
    int getAge() const { return age; }
    void setAge(int a) { age = a; }
    SexType getSex() const { return sex; }
    void setSex(SexType s) { sex = s; }
    bool isHCW() { return day_loc and (day_loc->getType() == HOSPITAL or day_loc->getType() == NURSINGHOME); }

    double getStartingNaturalEfficacy() const { return startingNaturalEfficacy; }
    void setStartingNaturalEfficacy(double v) { startingNaturalEfficacy = v; }
    int getNaturalImmunityDuration() const { return naturalImmunityDuration; }
    void setNaturalImmunityDuration(int v) { naturalImmunityDuration = v; }
    float getCrossProtectionProbability() const { return crossProtectionProbability; }
    void setCrossProtectionProbability(float v) { crossProtectionProbability = v; }
    int getVaccineImmunityDuration(size_t dose, StrainType strain) const { return _par->immunityDuration(immunityQuantile, dose, strain); }
    double getImmunityQuantile() const { return immunityQuantile; }
    void setImmunityQuantile(double v) { immunityQuantile = v; }
    bool getLongTermCare() { return long_term_care; }
    void setLongTermCare(bool b) { long_term_care = b; }
    ComorbidType hasComorbidity() { return comorbidity; }
    void setComorbidity(ComorbidType c) { comorbidity = c; }

    Location* getHomeLoc() { return home_loc; }
    void setHomeLoc(Location* loc) { home_loc = loc; }
    double getRiskiness() const { return home_loc->getRiskiness(); }
    Location* getDayLoc() { return day_loc; }
    void setDayLoc(Location* loc) { day_loc = loc; }
    Location* getHospital() const { return home_loc->getHospital(); }
    void addPatronizedLocation(Location* loc) { patronized_locs.push_back(loc); }
    std::vector<Location*> getPatronizedLocations() const { return patronized_locs; }
    void setPatronizedLocations(std::vector<Location*> pl) { patronized_locs = std::move(pl); }

    // ---- mirror of the device-side infection history --------------------------------------------------------------
    const std::vector<Infection*>& getInfectionHistory() const { fresh(); return infectionHistory; }
    int getNumNaturalInfections() const { fresh(); return (int) infectionHistory.size(); }
    bool hasBeenInfected() const { fresh(); return not infectionHistory.empty(); }
    Infection* getInfection(int infectionsago = 0) const { fresh(); return infectionHistory[infectionHistory.size() - 1 - infectionsago]; }
    Person* getInfectedBy(int ago = 0)      const { return getInfection(ago)->infectedBy; }
    Location* getInfectedPlace(int ago = 0) const { return getInfection(ago)->infectedPlace; }
    int getInfectedTime(int ago = 0)        const { return getInfection(ago)->infectedBegin; }
    int getInfectiousTime(int ago = 0)      const { return getInfection(ago)->infectiousBegin; }
    int getSymptomTime(int ago = 0)         const { return getInfection(ago)->symptomBegin; }
    int getSevereTime(int ago = 0)          const { return getInfection(ago)->severeBegin; }
    int getCriticalTime(int ago = 0)        const { return getInfection(ago)->criticalBegin; }
    int getHospitalizedTime(int ago = 0)    const { return getInfection(ago)->hospitalizedBegin; }
    int getIcuTime(int ago = 0)             const { return getInfection(ago)->icuBegin; }
    int getDeathTime(int ago = 0)           const { return getInfection(ago)->deathTime; }
    StrainType getStrain(int ago = 0)       const { return getInfection(ago)->strain; }
    double getRelInfectiousness(int ago = 0) const { return getInfection(ago)->relInfectiousness; }

    bool inHospital(int t)      const { return hasBeenInfected() and infectionHistory.back()->inHospital(t); }
    bool inIcu(int t)           const { return hasBeenInfected() and infectionHistory.back()->inIcu(t); }
    bool isDead(int t)          const { return hasBeenInfected() and infectionHistory.back()->isDead(t); }
    bool isAlive(int t)         const { return not isDead(t); }
    bool isNewlyInfected(int t) const { return hasBeenInfected() and t == infectionHistory.back()->infectedBegin; }
    bool isInfected(int t)      const { return hasBeenInfected() and infectionHistory.back()->isInfected(t); }
    bool isInfectious(int t)    const { return hasBeenInfected() and infectionHistory.back()->isInfectious(t); }
    bool isSymptomatic(int t)   const { return hasBeenInfected() and infectionHistory.back()->isSymptomatic(t); }
    bool isSevere(int t)        const { return hasBeenInfected() and infectionHistory.back()->isSevere(t); }
    bool isCritical(int t)      const { return hasBeenInfected() and infectionHistory.back()->isCritical(t); }
    bool isNewlyDead(int t)     const { return hasBeenInfected() and t == infectionHistory.back()->deathTime; }

    // ---- vaccination and quarantine: host copy written through to the device ------------------------------------
    int getNumVaccinations() const { return (int) vaccineHistory.size(); }
    const std::vector<int>& getVaccinationHistory() const { return vaccineHistory; }
    bool isVaccinated() const { return not vaccineHistory.empty(); }
    int daysSinceVaccination(int time) const { return time - vaccineHistory.back(); }
    bool isImmuneState(ImmuneStateType is) const;
    bool isSeroEligible() const;                       // Person.cpp:562-583
    bool isInfEligible(int today) const;               // Person.cpp:586-608
    void scheduleQuarantine(int start, int quarantineDuration);
    bool isQuarantining(int time) {
        return not quarantineHistory.empty() and time >= quarantineHistory.back().first and time < quarantineHistory.back().second;
    }

    // ---- infection entry points used by simulator.h (seed_epidemic :213-217) and Community::infect --------------
    Infection* infect(Community* community, Person* source, const Date* date, Location* sourceloc, StrainType strain = NUM_OF_STRAIN_TYPES,
                      bool check_susceptibility = true);
    Infection* infect(Community* community, const Date* date, StrainType strain) { return infect(community, nullptr, date, nullptr, strain); }

    static void setPar(const Parameters* par) { _par = par; }
    static void reset_ID_counter() { NEXT_ID = 0; }
    bool isSurveilledPerson() { return (int) id < _par->numSurveilledPeople; }
    double getBaselineRiskSevOutcome() { return _par->probSeriousOutcome.at(SEVERE)[hasComorbidity()][getAge()] * _par->pathogenicityByAge[getAge()]; }
    void revertState(const Date*) {}                   // the device snapshot restores everything (Community::load_from_cache)

  protected:
    size_t id;
    Location* home_loc;
    Location* day_loc;
    std::vector<Location*> patronized_locs;
    int age;
    SexType sex;
    bool long_term_care;
    ComorbidType comorbidity;
    mutable std::vector<Infection*> infectionHistory;
    int naturalImmunityDuration;
    double startingNaturalEfficacy;
    double immunityQuantile;
    float crossProtectionProbability;
    std::vector<int> vaccineHistory;
    std::vector<std::pair<int, int>> quarantineHistory;

    bool vaccinate(int time);                          // Person.cpp:611-647; reached through Vaccinee (Vac_Campaign.h:124)

    static void fresh() { if (_stale) refresh(); }     // pull new records of the device's infection log, once per change
    static void refresh();
    static bool _stale;
    static Community* _community;
    static const Parameters* _par;
    static size_t NEXT_ID;
};

struct PerPtrComp { bool operator()(const Person* a, const Person* b) const { return a->getID() < b->getID(); } };

#endif
