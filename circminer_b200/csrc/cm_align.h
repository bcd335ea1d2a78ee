#pragma once
#include "cm_types.h"
