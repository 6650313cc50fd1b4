// RoboGamePyB200.cpp -- the pybind11 translation unit a maintainer of the reference would build INSTEAD of
// src/RoboGamePy.cpp + src/GameSim.cpp (reference src/RoboGamePy.cpp:11-25, CMakeLists.txt:16): module
// `robo_game_py`, class `GameSim` {init, reset, undecided, run, __repr__} over the C ABI of librgs_b200.so
// (include/rgs_capi.h).  Needs neither Eigen nor pybind11/eigen.h.  INTEGRATION.md section B quotes this file;
// `make -C robo_game_sim_b200/pybind` builds it and tests/ run the drop-in checks through it as well as through
// the ctypes module the repository ships.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>

#include <stdexcept>
#include <string>

#include "rgs_capi.h"
namespace py = pybind11;
using arr = py::array_t<double, py::array::c_style | py::array::forcecast>;

namespace {

// pybind11/eigen.h converts any float-convertible 1-D length-n or (n,1) array into an Eigen vector; anything else is
// a TypeError at the call (reference behaviour)
const double *vec(const arr &a, py::ssize_t n, const char *what) {
    const bool ok = (a.ndim() == 1 && a.shape(0) == n) || (a.ndim() == 2 && a.shape(0) == n && a.shape(1) == 1);
    if (!ok) throw py::type_error(std::string(what) + ": expected " + std::to_string(n) + " doubles");
    return a.data();
}

struct GameSimB200 {
    rgs_single *h = nullptr;
    GameSimB200() {
        if (rgs_single_create(&h, 0) != RGS_OK) throw std::runtime_error(rgs_last_error());
    }
    ~GameSimB200() { rgs_single_destroy(h); }
    GameSimB200(const GameSimB200 &) = delete;
    GameSimB200 &operator=(const GameSimB200 &) = delete;

    arr reset(double dt, int winning_score, arr x0_ball, double noise, bool log, const std::string &logname, int seed,
              double tau_puck, double tau_player, double player_mass, double puck_mass) {
        const double *x0 = vec(x0_ball, 4, "x0_ball");
        arr out(22);
        const int rc = rgs_single_reset(h, dt, winning_score, x0, noise, log ? 1 : 0, logname.c_str(), seed, tau_puck, tau_player,
                                        player_mass, puck_mass, out.mutable_data());
        if (rc == RGS_EINVAL) throw py::value_error(rgs_last_error());
        if (rc != RGS_OK) throw std::runtime_error(rgs_last_error());
        return out;
    }
    bool undecided() {
        int32_t u = 0;
        if (rgs_single_undecided(h, &u) != RGS_OK) throw std::runtime_error(rgs_last_error());
        return u != 0;
    }
    arr run(arr a1, arr a2, arr b1, arr b2) {
        double u[8];
        const arr *v[4] = {&a1, &a2, &b1, &b2};
        for (int k = 0; k < 4; ++k) {
            const double *p = vec(*v[k], 2, "velocity command");
            u[2 * k] = p[0];
            u[2 * k + 1] = p[1];
        }
        arr out(22);
        {
            py::gil_scoped_release nogil;   // the reference holds the GIL; nothing here needs it
            if (rgs_single_run(h, u, out.mutable_data()) != RGS_OK) {
                py::gil_scoped_acquire gil;
                throw std::runtime_error(rgs_last_error());
            }
        }
        return out;
    }
};

}  // namespace

PYBIND11_MODULE(robo_game_py, m) {
    using namespace pybind11::literals;
    py::class_<GameSimB200>(m, "GameSim")
        .def(py::init())
        // the reference binds reset with 11 positional parameters and no defaults (src/RoboGamePy.cpp:17) although four
        // of its six scripts pass 6 (py/run_sim.py:43): the C++ defaults of src/GameSim.cpp:24-25 are declared here
        .def("reset", &GameSimB200::reset, "dt"_a, "winning_score"_a, "x0_ball"_a, "noise"_a, "log"_a, "logname"_a,
             "seed"_a = 0, "tau_puck"_a = 0.1, "tau_player"_a = 0.5, "player_mass"_a = 1.0, "puck_mass"_a = 0.5)
        .def("undecided", &GameSimB200::undecided)
        .def("run", &GameSimB200::run)
        .def("__repr__", [](const GameSimB200 &) { return "GameSim Object"; });
}
