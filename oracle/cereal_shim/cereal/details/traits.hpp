// Part of the oracle-only cereal stand-in; see cereal/cereal.hpp.
#pragma once
#include <cereal/cereal.hpp>
