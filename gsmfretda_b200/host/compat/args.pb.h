// compat/args.pb.h — the name protoc gives the generated header of args.proto; reference-side code
// that includes "args.pb.h" gets the hand-written codec (same message classes and accessors).
#pragma once
#include "../args_pb.hpp"
