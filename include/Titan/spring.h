// Titan/spring.h -- titan::Spring, the user-visible spring handle (drop-in for the reference's
// include/Titan/spring.h:16-75).  Host object; packed into the engine's per-spring arrays at start()/set()
// (tb_set_spring_params); get() brings back only _rest, the one field the device mutates (src/spring.cu:12-14).
#ifndef TITAN_B200_SPRING_H
#define TITAN_B200_SPRING_H

#include <cstdint>

#include "mass.h"
#include "vec.h"

namespace titan {

enum SpringType { PASSIVE_SOFT, PASSIVE_STIFF, ACTIVE_CONTRACT_THEN_EXPAND, ACTIVE_EXPAND_THEN_CONTRACT,
                  ACTUATED_EXPAND, ACTUATED_CONTRACT };

class Spring {
public:
    Spring() : _left(nullptr), _right(nullptr), _k(10000.0), _rest(1.0), _type(PASSIVE_SOFT), _omega(0.0),
               _damping(0.0), _l_max(0.0), _l_min(0.0), _rate(0.0), slot(-1) {}
    Spring(Mass *left, Mass *right) : Spring() { _left = left; _right = right; defaultLength(); }
    Spring(Mass *left, Mass *right, double k, double rest_length) : Spring() { _left = left; _right = right; _k = k; _rest = rest_length; }
    Spring(Mass *left, Mass *right, double k, double rest_length, SpringType type, double omega, double max_length,
           double min_length, double expansion_rate) : Spring(left, right, k, rest_length) {
        _type = type; _omega = omega; _l_max = max_length; _l_min = min_length; _rate = expansion_rate;
    }

    void setRestLength(double rest_length) { _rest = rest_length; }
    void defaultLength();  // |left.pos - right.pos|
    void changeType(SpringType type, double omega) { _type = type; _omega = omega; }
    void addDamping(double constant) { _damping = constant; }
    void setLeft(Mass *left);
    void setRight(Mass *right);
    void setMasses(Mass *left, Mass *right) { _left = left; _right = right; }

    Mass *_left;
    Mass *_right;
    double _k;        // N/m
    double _rest;     // m
    SpringType _type;
    double _omega;    // actuation frequency (ACTIVE_* types)
    double _damping;
    double _l_max;    // actuator limits and rate (ACTUATED_* types); zero-initialised here
    double _l_min;
    double _rate;

private:
    int32_t slot;     // index in the engine's arrays, -1 while not uploaded
    friend class Simulation;
    friend class Container;
};

}  // namespace titan

#endif
