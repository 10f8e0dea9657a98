// <glm/gtc/matrix_transform.hpp> stand-in: everything the reference's tests use is in ../glm.hpp
#pragma once
#include "../glm.hpp"
