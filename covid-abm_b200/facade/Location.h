// covid-abm_b200/facade/Location.h -- drop-in for the reference's Location.h (Location.h:11-110).
//
// Host-side proxy of one location. All per-day state of a location (infectious pressure, visitors, admitted
// patients) lives on the GPU behind include/cabm.h; this object only keeps what the reference's callers read:
// id, type, flags, coordinates, the fixed roster (residents / staff / students), household neighbours and the
// associated hospital. Included after Parameters.h, like the original (it uses LocationType et al.).
#ifndef CABM_FACADE_LOCATION_H
#define CABM_FACADE_LOCATION_H

#include <set>
#include <utility>
#include <vector>

class Person;

class Location {
  public:
    struct LocPtrComp { bool operator()(const Location* a, const Location* b) const { return a->getID() < b->getID(); } };

    Location() : _ID((int) NEXT_ID++), _essential(true), _riskiness(0.0f), _public_transmission_risk(NO_PUBLIC_TRANSMISSION),
                 _type(NUM_OF_LOCATION_TYPES), _hospital(nullptr), _coord(0.0, 0.0), _xPixel(0.0), _yPixel(0.0) {}
    virtual ~Location() {}

    int getID() const { return _ID; }
    static void reset_ID_counter() { NEXT_ID = 0; }

    LocationType getType() const { return _type; }
    void setType(LocationType t) { _type = t; }
    void setEssential(bool e) { _essential = e; }
    bool isEssential() const { return _essential; }
    bool isNonEssential() const { return !_essential; }

    // the fixed roster; daily visitors and admitted patients are device state and are not mirrored here
    int getNumPeople() const { return (int) _person.size(); }
    std::vector<Person*> getPeople() { return _person; }
    const std::vector<Person*>& people() const { return _person; }
    Person* getPerson(int idx) { return _person[idx]; }
    void setPeople(std::vector<Person*> p) { _person = std::move(p); }
    void addPerson(Person* p) { _person.push_back(p); }
    bool removePerson(Person* p) {
        for (size_t i = 0; i < _person.size(); ++i) if (_person[i] == p) { _person[i] = _person.back(); _person.pop_back(); return true; }
        return false;
    }
    int getNumVisitors() const { return 0; }
    std::vector<Person*> getVisitors() { return {}; }
    void clearVisitors() {}

    std::set<Location*, LocPtrComp> getNeighbors() { return _neighbors; }
    const std::set<Location*, LocPtrComp>& neighbors() const { return _neighbors; }
    int getNumNeighbors() const { return (int) _neighbors.size(); }
    void setNeighbors(std::set<Location*, LocPtrComp> n) { _neighbors = std::move(n); }
    void addNeighbor(Location* loc) { _neighbors.insert(loc); }

    Location* getHospital() const { return _hospital; }
    void setHospital(Location* h) { _hospital = h; }

    void setCoordinates(std::pair<double, double> c) { _coord = c; }
    std::pair<double, double> getCoordinates() { return _coord; }
    double getX() const { return _coord.first; }
    void setX(double x) { _coord.first = x; }
    double getY() const { return _coord.second; }
    void setY(double y) { _coord.second = y; }
    void setPixel(double xP, double yP) { _xPixel = xP; _yPixel = yP; }
    std::pair<double, double> getPixel() const { return { _xPixel, _yPixel }; }
    double getXPixel() const { return _xPixel; }
    double getYPixel() const { return _yPixel; }

    float getRiskiness() const { return _riskiness; }
    void setRiskiness(float r) { _riskiness = r; }
    PublicTransmissionType getPublicTransmissionRisk() const { return _public_transmission_risk; }
    void setPublicTransmissionRisk(PublicTransmissionType pt) { _public_transmission_risk = pt; }

    bool operator==(const Location* other) const { return _ID == other->_ID; }
    void revertState() {}

  protected:
    int _ID;
    bool _essential;
    float _riskiness;
    PublicTransmissionType _public_transmission_risk;
    LocationType _type;
    std::vector<Person*> _person;
    std::set<Location*, LocPtrComp> _neighbors;
    Location* _hospital;
    std::pair<double, double> _coord;
    double _xPixel, _yPixel;
    static size_t NEXT_ID;
};

#endif
