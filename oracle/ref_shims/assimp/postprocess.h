#pragma once
#include "cimport.h"
