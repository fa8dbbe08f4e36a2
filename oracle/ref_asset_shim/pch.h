/* pch.h - stand-in for the reference's precompiled header (/root/reference/src/pch.h pulls in Windows, Vulkan and SDL
 * headers that are absent here) so that the reference's own src/Asset/Asset.cpp and src/Logger/Logger.cpp compile
 * where they lie with g++.  Test infrastructure for oracle/_ref/libasset_ref.so; holds no reference code: only the
 * standard headers those two files use and the two MSVC-isms they rely on. */
#pragma once
#include <array>
#include <cassert>
#include <cerrno>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <filesystem>
#include <format>
#include <fstream>
#include <iostream>
#include <memory>
#include <string>
#include <string_view>
#include <vector>

typedef int errno_t;
static inline errno_t fopen_s(FILE** stream, const char* path, const char* mode) {
    *stream = std::fopen(path, mode);
    return *stream ? 0 : (errno ? errno : 1);
}
