// Minimal stand-in so that the reference headers compile without Boost (oracle/_ref build only).
#pragma once
#define BOOST_VERSION 105400
