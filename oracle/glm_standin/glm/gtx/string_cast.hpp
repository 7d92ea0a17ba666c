// ORACLE / TEST INFRASTRUCTURE ONLY.  The reference includes <glm/gtx/string_cast.hpp>
// (billiard_search.cpp:8) but uses glm::to_string only inside DEBUG(...) statements, which
// are compiled out under -DNDEBUG (billiard_debug.hpp:8-21).  Nothing to provide.
#pragma once
#include <glm/glm.hpp>
