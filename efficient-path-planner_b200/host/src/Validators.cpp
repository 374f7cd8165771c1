// Validators.cpp — StateValidator / MotionValidator bodies (reference: src/StateValidator.cpp:7-13,
// src/MotionValidator.cpp:8-23).
#include <iostream>

#include "epp/MotionValidator.h"
#include "epp/StateValidator.h"

namespace ob = ompl::base;

bool StateValidator::isValid(const ompl::base::State *state) const
{
    const ob::RealVectorStateSpace::StateType *state3D = state->as<ob::RealVectorStateSpace::StateType>();
    const epp::Vector3d pos(state3D->values[0], state3D->values[1], state3D->values[2]);
    return world->checkPointValidity(pos, canPassGate);
}

void StateValidator::isValidBatch(const ompl::base::State *const *states, std::size_t n, uint32_t *maskWords) const
{
    std::vector<double> xyz(3 * n);
    for (std::size_t i = 0; i < n; ++i) {
        const double *v = states[i]->as<ob::RealVectorStateSpace::StateType>()->values;
        xyz[3 * i] = v[0];
        xyz[3 * i + 1] = v[1];
        xyz[3 * i + 2] = v[2];
    }
    world->checkPointValidityBatch(xyz.data(), n, canPassGate, maskWords);
}

void StateValidator::isValidBatch(const double *xyz, std::size_t n, uint32_t *maskWords) const
{
    world->checkPointValidityBatch(xyz, n, canPassGate, maskWords);
}

std::vector<bool> StateValidator::isValidBatch(const std::vector<const ompl::base::State *> &states) const
{
    std::vector<uint32_t> words(World::maskWords(states.size()));
    isValidBatch(states.data(), states.size(), words.data());
    std::vector<bool> out(states.size());
    for (std::size_t i = 0; i < states.size(); ++i) out[i] = (words[i >> 5] >> (i & 31)) & 1u;
    return out;
}

bool MotionValidator::checkMotion(const ob::State *s1, const ob::State *s2) const
{
    const ob::RealVectorStateSpace::StateType *state1 = s1->as<ob::RealVectorStateSpace::StateType>();
    const ob::RealVectorStateSpace::StateType *state2 = s2->as<ob::RealVectorStateSpace::StateType>();
    const epp::Vector3d start(state1->values[0], state1->values[1], state1->values[2]);
    const epp::Vector3d end(state2->values[0], state2->values[1], state2->values[2]);
    return world->checkRayValid(start, end, canPassGate);
}

bool MotionValidator::checkMotion(const ob::State *, const ob::State *, std::pair<ob::State *, double> &) const
{
    std::cerr << "Not implemented error" << std::endl;
    return true;
}

void MotionValidator::checkMotionBatch(const ob::State *const *s1, const ob::State *const *s2, std::size_t n,
                                       uint32_t *maskWords) const
{
    std::vector<double> a(3 * n), b(3 * n);
    for (std::size_t i = 0; i < n; ++i) {
        const double *va = s1[i]->as<ob::RealVectorStateSpace::StateType>()->values;
        const double *vb = s2[i]->as<ob::RealVectorStateSpace::StateType>()->values;
        for (int d = 0; d < 3; ++d) {
            a[3 * i + d] = va[d];
            b[3 * i + d] = vb[d];
        }
    }
    world->checkRayValidBatch(a.data(), b.data(), n, canPassGate, maskWords);
}

void MotionValidator::checkMotionBatch(const double *starts, const double *ends, std::size_t n, uint32_t *maskWords) const
{
    world->checkRayValidBatch(starts, ends, n, canPassGate, maskWords);
}
