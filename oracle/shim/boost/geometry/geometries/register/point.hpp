#pragma once
#include "../../../geometry.hpp"
