// solver_b200.h -- the public interface of the reference's Solver<D> (Microchannel3D/*/solver.h:10-86) on top of
// liblbm_b200.so, so that a main.cpp written against the reference compiles against this header with
// `Solver<Descriptor>` replaced by `Solver`: same method names, same arguments, same files written.
//   * LBM_GPUS=N in the environment (default 1) spreads the grid over N GPUs of this machine as N z-slabs with the
//     reference's own slab rule (solver.hpp:24-30): one host thread and one context per GPU (devices LBM_DEVICE ..
//     LBM_DEVICE+N-1), halos exchanged inside the library over NCCL (lbm_comm_attach), put* / getWhole* / the writers
//     work on the whole grid exactly as the reference's rank 0 does after its gather (solver.hpp:121-277) -- the
//     replacement of `mpiexec -n N` (binary.pbs:58) on a single multi-GPU box.  Results do not depend on N.
//   * every .vts is the file VTK 5's vtkXMLStructuredGridWriter would write in raw appended mode
//   * additions with no reference counterpart are marked "extension"
#ifndef LBM_SOLVER_B200_H
#define LBM_SOLVER_B200_H
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <stdexcept>
#include <thread>
#include <unistd.h>
#include "host_common.h"

// params_list.h: name -> value(s); the hosts only ever add one number per name
class ParamsList {
public:
    template <typename T>
    void add(const std::string &name, T v) { values_[name] = (double)v; }
    bool has(const std::string &name) const { return values_.count(name) != 0; }
    double get(const std::string &name) const
    {
        std::map<std::string, double>::const_iterator it = values_.find(name);
        if (it == values_.end()) throw std::runtime_error("Parameter " + name + " was not found.");
        return it->second;
    }

private:
    std::map<std::string, double> values_;
};

class Solver {
public:
    // solver.hpp:17-47 (Solver) + lattice.hpp constructor: sizes and constants from the list
    explicit Solver(const ParamsList &p) : NX((int)p.get("NX")), NY((int)p.get("NY")), NZ((int)p.get("NZ"))
    {
        lbm_config cfg;
        lbm_config_default(&cfg, LBM_BINARY_D3Q19);
        cfg.nx = NX; cfg.ny = NY; cfg.nz = NZ;
        cfg.aconst = p.get("aconst"); cfg.kconst = p.get("kconst"); cfg.gammaconst = p.get("gammaconst");
        cfg.wall_gradient = p.get("phase_gradient"); cfg.rho_wall = p.get("rho_wall");
        cfg.tau_phi = p.get("tau_phi"); cfg.tau_liq = p.get("tau_liq"); cfg.tau_gas = p.get("tau_gas");
        cfg.force[0] = p.get("force_x"); cfg.force[1] = p.get("force_y"); cfg.force[2] = p.get("force_z");
        // LBM_KERNEL_VARIANT: lbm_config.kernel_variant (default 0); e.g. 2 runs the step in place on one population array
        // (half the population memory; one GPU only)
        if (const char *v = std::getenv("LBM_KERNEL_VARIANT")) cfg.kernel_variant = std::atoi(v);
        const char *e = std::getenv("LBM_GPUS");
        const int ngpu = e ? std::atoi(e) : 1;
        if (ngpu < 1 || NZ / ngpu < 2) throw std::runtime_error("LBM_GPUS: need at least 2 z planes per GPU");
        slabs_.resize(ngpu);
        for (int r = 0; r < ngpu; r++) {
            cfg.device = device_from_env() + r;
            cfg.slab_rank = r; cfg.slab_count = ngpu;
            LBM_CHECK(0, lbm_create(&cfg, &slabs_[r].ctx));
            int z0 = 0, nzl = 0;
            LBM_CHECK(slabs_[r].ctx, lbm_slab_extent(slabs_[r].ctx, &z0, &nzl));
            slabs_[r].first = (size_t)z0 * NX * NY;
            slabs_[r].count = (size_t)nzl * NX * NY;
        }
        if (ngpu > 1) {
            // one communicator over the slabs: every rank has to join at the same time, hence one thread each
            char id[LBM_COMM_ID_BYTES];
            LBM_CHECK(0, lbm_comm_unique_id(id));
            // NCCL prints a version banner on stdout at the first communicator when NCCL_DEBUG=VERSION is set; the
            // program's stdout has to stay the reference's, so file descriptor 1 points at stderr meanwhile
            std::cout.flush();
            std::fflush(stdout);
            const int saved = dup(1);
            dup2(2, 1);
            each_slab([&](Slab &s) { LBM_CHECK(s.ctx, lbm_comm_attach(s.ctx, id)); });
            std::fflush(stdout);
            dup2(saved, 1);
            close(saved);
        }
        n_ = (size_t)NX * NY * NZ;
        rho_.assign(n_, 0.0); phi_.assign(n_, 0.0); ux_.assign(n_, 0.0); uy_.assign(n_, 0.0); uz_.assign(n_, 0.0);
        density.assign(n_, 0.0); phase.assign(n_, 0.0); velocity.assign(3 * n_, 0.0);
    }
    ~Solver()
    {
        for (size_t r = 0; r < slabs_.size(); r++) lbm_destroy(slabs_[r].ctx);
    }

    // solver.hpp:49-55 probes one node of the last gathered phase array; this probes every population on the device
    bool checkNAN()
    {
        bool any = false;
        for (size_t r = 0; r < slabs_.size(); r++) {
            int has_nan = 0;
            LBM_CHECK(slabs_[r].ctx, lbm_diagnostics(slabs_[r].ctx, 0, 0, 0, &has_nan));
            any = any || has_nan != 0;
        }
        return any;
    }

    // solver.hpp:58-118 / lattice.hpp putDensity...: inclusive index boxes into the fields init() starts from
    void putDensity(int x0, int y0, int z0, int x1, int y1, int z1, double dense) { put(rho_, x0, y0, z0, x1, y1, z1, dense); }
    void putDensity(double dense) { put(rho_, 0, 0, 0, NX - 1, NY - 1, NZ - 1, dense); }
    void putPhase(int x0, int y0, int z0, int x1, int y1, int z1, double v) { put(phi_, x0, y0, z0, x1, y1, z1, v); }
    void putPhase(double v) { put(phi_, 0, 0, 0, NX - 1, NY - 1, NZ - 1, v); }
    void putVelocity(int x0, int y0, int z0, int x1, int y1, int z1, double *u)
    {
        put(ux_, x0, y0, z0, x1, y1, z1, u[0]); put(uy_, x0, y0, z0, x1, y1, z1, u[1]); put(uz_, x0, y0, z0, x1, y1, z1, u[2]);
    }
    void putVelocity(double *u) { putVelocity(0, 0, 0, NX - 1, NY - 1, NZ - 1, u); }

    // solver.hpp:871-879
    void init()
    {
        each_slab([&](Slab &s) {
            LBM_CHECK(s.ctx, lbm_set_field(s.ctx, LBM_RHO, rho_.data() + s.first, s.count));
            LBM_CHECK(s.ctx, lbm_set_field(s.ctx, LBM_PHASE, phi_.data() + s.first, s.count));
            LBM_CHECK(s.ctx, lbm_set_field(s.ctx, LBM_UX, ux_.data() + s.first, s.count));
            LBM_CHECK(s.ctx, lbm_set_field(s.ctx, LBM_UY, uy_.data() + s.first, s.count));
            LBM_CHECK(s.ctx, lbm_set_field(s.ctx, LBM_UZ, uz_.data() + s.first, s.count));
            LBM_CHECK(s.ctx, lbm_init(s.ctx));
        });
    }
    // solver.hpp:881-891; collide_stream(n) (extension) queues n steps at once and only the last one leaves the
    // macroscopic fields behind, which is all the writers can see in the reference too
    void collide_stream() { collide_stream(1); }
    void collide_stream(int nsteps)
    {
        each_slab([&](Slab &s) { LBM_CHECK(s.ctx, lbm_step(s.ctx, nsteps)); });
    }

    // solver.hpp:121-277: every slab's part lands at its place in the public arrays density, phase, velocity
    // (interleaved 3*i+c) -- the gather to rank 0 of the reference
    void getWholeDensity() { get_whole(LBM_RHO, density); }
    void getWholePhase() { get_whole(LBM_PHASE, phase); }
    void getWholeVelocity()
    {
        get_whole(LBM_UX, ux_); get_whole(LBM_UY, uy_); get_whole(LBM_UZ, uz_);
        for (size_t i = 0; i < n_; i++) { velocity[3 * i] = ux_[i]; velocity[3 * i + 1] = uy_[i]; velocity[3 * i + 2] = uz_[i]; }
    }

    // ---- .vts writers: solver.hpp:280-398 (one array each) and :523-576 (Velocity, Phase, Density in that order) ----
    void writeWholeDensity(std::string name) { getWholeDensity(); VtsArray a[1] = {{"Density", 1, density.data()}}; write_vts(name, a, 1); }
    void writeWholePhase(std::string name) { getWholePhase(); VtsArray a[1] = {{"Phase", 1, phase.data()}}; write_vts(name, a, 1); }
    void writeWholeVelocity(std::string name) { getWholeVelocity(); VtsArray a[1] = {{"Velocity", 3, velocity.data()}}; write_vts(name, a, 1); }
    void writeWholeDensityPhaseVelocity(std::string name)
    {
        getWholeDensity(); getWholePhase(); getWholeVelocity();
        VtsArray a[3] = {{"Velocity", 3, velocity.data()}, {"Phase", 1, phase.data()}, {"Density", 1, density.data()}};
        write_vts(name, a, 3);
    }

    // ---- text slice writers, solver.hpp:578-755: "i j value(s)" per line, default stream precision unless stated ----
    void writeTextXZVelocity(std::string name)
    {
        getWholeVelocity();
        std::ofstream fout((name + ".dat").c_str());
        for (int z = 0; z < NZ; z++)
            for (int x = 0; x < NX; x++) vel_line(fout << x << " " << z << " ", at(x, NY / 2, z));
    }
    void writeTextXYVelocity(std::string name)
    {
        getWholeVelocity();
        std::ofstream fout((name + ".dat").c_str());
        for (int y = 0; y < NY; y++)
            for (int x = 0; x < NX; x++) vel_line(fout << x << " " << y << " ", at(x, y, NZ / 2));
    }
    void writeTextYZVelocity(std::string name)
    {
        getWholeVelocity();
        std::ofstream fout((name + ".dat").c_str());
        for (int z = 0; z < NZ; z++)
            for (int y = 0; y < NY; y++) vel_line(fout << y << " " << z << " ", at(NX / 2, y, z));
    }
    void writeTextWholeVelocity(std::string name)          // precision 12
    {
        getWholeVelocity();
        std::ofstream fout((name + ".dat").c_str());
        fout.precision(12);
        for (int z = 0; z < NZ; z++)
            for (int y = 0; y < NY; y++)
                for (int x = 0; x < NX; x++) vel_line(fout << x << " " << y << " " << z << " ", at(x, y, z));
    }
    void writeTextWholePhase(std::string name) { getWholePhase(); text_whole(name, phase, 0); }
    void writeTextWholeDensity(std::string name) { getWholeDensity(); text_whole(name, density, 10); }   // precision 10
    void writeTextXYPhase(std::string name) { getWholePhase(); text_xy(name, phase); }
    void writeTextXZPhase(std::string name) { getWholePhase(); text_xz(name, phase); }
    void writeTextYZPhase(std::string name) { getWholePhase(); text_yz(name, phase); }
    void writeTextXYDensity(std::string name) { getWholeDensity(); text_xy(name, density); }
    void writeTextXZDensity(std::string name) { getWholeDensity(); text_xz(name, density); }
    void writeTextYZDensity(std::string name) { getWholeDensity(); text_yz(name, density); }

    // solver.hpp:758-787: legacy ASCII STRUCTURED_GRID with the phase field
    void writeVTKWholePhase(std::string name)
    {
        getWholePhase();
        std::ofstream fout((name + ".vtk").c_str());
        fout << "# vtk DataFile Version 3.0\n";
        fout << "Binary liquid phase field vtk representation\n";
        fout << "ASCII\n\n";
        fout << "DATASET STRUCTURED_GRID\n";
        fout << "DIMENSIONS " << NX << " " << NY << " " << NZ << "\n";
        fout << "POINTS " << NX * NY * NZ << " double\n";
        for (int z = 0; z < NZ; z++)
            for (int y = 0; y < NY; y++)
                for (int x = 0; x < NX; x++) fout << x << " " << y << " " << z << "\n";
        fout << "\n";
        fout << "POINT_DATA " << NX * NY * NZ << "\n";
        fout << "SCALARS phase double\n";
        fout << "LOOKUP_TABLE phase_table\n";
        for (size_t i = 0; i < n_; i++) fout << phase[i] << "\n";
    }
    // per-rank files of solver.hpp:790-818; one rank here, so "_process0"
    void writeLocalDensity(std::string name) { getWholeDensity(); VtsArray a[1] = {{"Density", 1, density.data()}}; write_vts(name + "_process0", a, 1); }
    void writeLocalPhase(std::string name) { getWholePhase(); VtsArray a[1] = {{"Phase", 1, phase.data()}}; write_vts(name + "_process0", a, 1); }
    void writeLocalTextDensity(std::string name) { getWholeDensity(); text_whole(name + "_process0", density, 10); }
    void writeLocalTextPhase(std::string name) { getWholePhase(); text_whole(name + "_process0", phase, 0); }

    // solver.hpp:400-520: macroscopic restart -- Density, Phase and Velocity of a .vts (this writer's or any
    // raw-appended UInt64-header file with those arrays) become the fields the next init() starts from.  Lossy by
    // design, as in the reference: populations are rebuilt as equilibria.
    void load_file(std::string name)
    {
        std::vector<double> v;
        read_vts_array(name + ".vts", "Density", 1, rho_);
        read_vts_array(name + ".vts", "Phase", 1, phi_);
        read_vts_array(name + ".vts", "Velocity", 3, v);
        for (size_t i = 0; i < n_; i++) { ux_[i] = v[3 * i]; uy_[i] = v[3 * i + 1]; uz_[i] = v[3 * i + 2]; }
    }

    // extension: exact restart.  The file holds both population sets as raw doubles [node*19+k] behind
    // a small header; a run continued from it is bit-identical to the uninterrupted run.
    void savePopulations(std::string name, int time_counter)
    {
        std::vector<double> pop(n_ * 19);
        std::ofstream out((name + ".pop").c_str(), std::ios::binary);
        const int64_t head[6] = {0x504f50324c424dLL, NX, NY, NZ, 19, time_counter};
        out.write((const char *)head, sizeof head);
        for (int set = 0; set < 2; set++) {
            each_slab([&](Slab &s) { LBM_CHECK(s.ctx, lbm_get_populations(s.ctx, set, pop.data() + s.first * 19, s.count * 19)); });
            out.write((const char *)pop.data(), pop.size() * sizeof(double));
        }
    }
    int loadPopulations(std::string name)                  // returns the time_counter stored with the dump
    {
        std::ifstream in((name + ".pop").c_str(), std::ios::binary);
        int64_t head[6];
        in.read((char *)head, sizeof head);
        if (!in || head[0] != 0x504f50324c424dLL || head[1] != NX || head[2] != NY || head[3] != NZ || head[4] != 19)
            throw std::runtime_error("loadPopulations: " + name + ".pop does not match this solver");
        std::vector<double> pop(n_ * 19);
        // the wall and mirror nodes of the reference hold junk that never enters the update; init() first gives
        // every auxiliary array a defined state, then the populations are replaced
        init();
        for (int set = 0; set < 2; set++) {
            in.read((char *)pop.data(), pop.size() * sizeof(double));
            if (!in) throw std::runtime_error("loadPopulations: file too short");
            // all slabs at once: setting populations also refreshes the ghost planes from the neighbours
            each_slab([&](Slab &s) { LBM_CHECK(s.ctx, lbm_set_populations(s.ctx, set, pop.data() + s.first * 19, s.count * 19)); });
        }
        return (int)head[5];
    }

    lbm_ctx *context(int slab = 0) { return slabs_[slab].ctx; }
    int gpus() const { return (int)slabs_.size(); }
    const int NX, NY, NZ;
    std::vector<double> density, phase, velocity;          // public in spirit: what getWhole* fill (solver.h:82-84)

private:
    struct VtsArray { const char *name; int ncomp; const double *data; };
    struct Slab { lbm_ctx *ctx; size_t first, count; };   // context of one GPU; its nodes are [first, first+count) of the grid
    std::vector<Slab> slabs_;
    size_t n_;

    // runs fn on every slab: directly with one GPU, one host thread per GPU otherwise (calls that exchange halos
    // block until the neighbours have made the same call)
    template <class Fn>
    void each_slab(Fn fn)
    {
        if (slabs_.size() == 1) { fn(slabs_[0]); return; }
        std::vector<std::thread> th;
        for (size_t r = 0; r < slabs_.size(); r++) th.push_back(std::thread([this, r, &fn] { fn(slabs_[r]); }));
        for (size_t r = 0; r < th.size(); r++) th[r].join();
    }
    void get_whole(int field, std::vector<double> &dst)
    {
        each_slab([&](Slab &s) { LBM_CHECK(s.ctx, lbm_get_field(s.ctx, field, dst.data() + s.first, s.count)); });
    }
    std::vector<double> rho_, phi_, ux_, uy_, uz_;         // what put*/load_file fill and init() uploads

    size_t at(int x, int y, int z) const { return ((size_t)z * NY + y) * NX + x; }
    void put(std::vector<double> &a, int x0, int y0, int z0, int x1, int y1, int z1, double v)
    {
        if (z1 > NZ - 1) z1 = NZ - 1;
        if (z0 < 0) z0 = 0;
        for (int z = z0; z <= z1; z++)
            for (int y = y0; y <= y1; y++)
                for (int x = x0; x <= x1; x++) a[at(x, y, z)] = v;
    }
    void vel_line(std::ostream &o, size_t i) const { o << velocity[3 * i] << " " << velocity[3 * i + 1] << " " << velocity[3 * i + 2] << "\n"; }
    void text_whole(const std::string &name, const std::vector<double> &a, int precision)
    {
        std::ofstream fout((name + ".dat").c_str());
        if (precision) fout.precision(precision);
        for (int z = 0; z < NZ; z++)
            for (int y = 0; y < NY; y++)
                for (int x = 0; x < NX; x++) fout << x << " " << y << " " << z << " " << a[at(x, y, z)] << "\n";
    }
    void text_xy(const std::string &name, const std::vector<double> &a)
    {
        std::ofstream fout((name + ".dat").c_str());
        for (int y = 0; y < NY; y++)
            for (int x = 0; x < NX; x++) fout << x << " " << y << " " << a[at(x, y, NZ / 2)] << "\n";
    }
    void text_xz(const std::string &name, const std::vector<double> &a)
    {
        std::ofstream fout((name + ".dat").c_str());
        for (int z = 0; z < NZ; z++)
            for (int x = 0; x < NX; x++) fout << x << " " << z << " " << a[at(x, NY / 2, z)] << "\n";
    }
    void text_yz(const std::string &name, const std::vector<double> &a)
    {
        std::ofstream fout((name + ".dat").c_str());
        for (int z = 0; z < NZ; z++)
            for (int y = 0; y < NY; y++) fout << y << " " << z << " " << a[at(NX / 2, y, z)] << "\n";
    }

    // XML StructuredGrid, raw appended data with UInt64 block headers: the point-data arrays in the order given, then
    // the Float32 points
    void write_vts(const std::string &name, const VtsArray *arr, int narr)
    {
        std::ofstream out((name + ".vts").c_str(), std::ios::binary);
        uint64_t off = 0;
        out << "<?xml version=\"1.0\"?>\n<VTKFile type=\"StructuredGrid\" version=\"1.0\" byte_order=\"LittleEndian\" header_type=\"UInt64\">\n";
        out << "  <StructuredGrid WholeExtent=\"0 " << NX - 1 << " 0 " << NY - 1 << " 0 " << NZ - 1 << "\">\n";
        out << "    <Piece Extent=\"0 " << NX - 1 << " 0 " << NY - 1 << " 0 " << NZ - 1 << "\">\n      <PointData>\n";
        for (int a = 0; a < narr; a++) {
            out << "        <DataArray type=\"Float64\" Name=\"" << arr[a].name << "\"";
            if (arr[a].ncomp != 1) out << " NumberOfComponents=\"" << arr[a].ncomp << "\"";
            out << " format=\"appended\" offset=\"" << off << "\"/>\n";
            off += 8 + (uint64_t)n_ * arr[a].ncomp * 8;
        }
        out << "      </PointData>\n      <CellData/>\n      <Points>\n";
        out << "        <DataArray type=\"Float32\" Name=\"Points\" NumberOfComponents=\"3\" format=\"appended\" offset=\"" << off << "\"/>\n";
        out << "      </Points>\n    </Piece>\n  </StructuredGrid>\n  <AppendedData encoding=\"raw\">\n   _";
        for (int a = 0; a < narr; a++) {
            const uint64_t bytes = (uint64_t)n_ * arr[a].ncomp * 8;
            out.write((const char *)&bytes, 8);
            out.write((const char *)arr[a].data, bytes);
        }
        std::vector<float> pts(n_ * 3);
        size_t i = 0;
        for (int z = 0; z < NZ; z++)
            for (int y = 0; y < NY; y++)
                for (int x = 0; x < NX; x++) { pts[i++] = (float)x; pts[i++] = (float)y; pts[i++] = (float)z; }
        const uint64_t bytes_p = (uint64_t)n_ * 3 * 4;
        out.write((const char *)&bytes_p, 8);
        out.write((const char *)pts.data(), bytes_p);
        out << "\n  </AppendedData>\n</VTKFile>\n";
    }

    // finds <DataArray ... Name="name" ... offset="o"/> and reads its block out of the raw appended section
    void read_vts_array(const std::string &file, const char *name, int ncomp, std::vector<double> &dst)
    {
        std::ifstream in(file.c_str(), std::ios::binary);
        if (!in) throw std::runtime_error("load_file: cannot open " + file);
        std::string all((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
        const size_t tag = all.find(std::string("Name=\"") + name + "\"");
        const size_t app = all.find("<AppendedData");
        if (tag == std::string::npos || app == std::string::npos || tag > app) throw std::runtime_error(std::string("load_file: no array ") + name);
        const size_t end = all.find("/>", tag), o = all.find("offset=\"", tag);
        if (o == std::string::npos || o > end) throw std::runtime_error("load_file: only appended data is supported");
        const uint64_t off = std::strtoull(all.c_str() + o + 8, 0, 10);
        const size_t base = all.find('_', app) + 1 + off;
        uint64_t bytes = 0;
        if (base + 8 > all.size()) throw std::runtime_error("load_file: truncated file");
        std::memcpy(&bytes, all.data() + base, 8);
        if (bytes != (uint64_t)n_ * ncomp * 8 || base + 8 + bytes > all.size()) throw std::runtime_error(std::string("load_file: wrong size of ") + name);
        dst.resize(n_ * ncomp);
        std::memcpy(dst.data(), all.data() + base + 8, bytes);
    }
};
#endif
