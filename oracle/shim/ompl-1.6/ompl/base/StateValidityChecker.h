#pragma once
#include "../../../../../efficient-path-planner_b200/host/ompl_shim/ompl/base/StateValidityChecker.h"
