"""the Parthenon-side adapter is real code. Parthenon is absent from this image, so
integration/jaybenne_adapter.cpp — jaybenne::Initialize / RadiationStep / InitializeRadiation / the public tasks /
PhotonReflectBC on top of the C ABI — is type-checked against declaration-only stand-ins of the Parthenon and
singularity types it touches (integration/mock/), and checked to define every symbol the reference's
src/jaybenne/jaybenne.hpp:50-76 declares and to read every <jaybenne> key of src/jaybenne/jaybenne.cpp:163-223."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ADAPTER = os.path.join(ROOT, "integration", "jaybenne_adapter.cpp")


def test_adapter_type_checks_against_the_mock_parthenon():
    r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-Wextra", "-Werror",
                        "-I" + os.path.join(ROOT, "integration", "mock"), "-I" + os.path.join(ROOT, "include"), ADAPTER],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_adapter_defines_the_reference_package_interface():
    src = open(ADAPTER).read()
    for sym in ("std::shared_ptr<StateDescriptor> Initialize(ParameterInput *pin, Opacity &opacity, Scattering &scattering, EOS &eos)",
                "TaskCollection RadiationStep(Mesh *pmesh, const Real t_start, const Real dt)",
                "void InitializeRadiation(MeshBlockData<Real> *mbd, const bool is_thermal)",
                "Real EstimateTimestepMesh(MeshData<Real> *md)",
                "TaskStatus TransportPhotons(MeshData<Real> *md, const Real t_start, const Real dt)",
                "TaskStatus TransportPhotons_DDMC(MeshData<Real> *md, const Real t_start, const Real dt)",
                "TaskStatus SampleDDMCBlockFace(MeshData<Real> *md)",
                "TaskStatus CheckCompletion(MeshData<Real> *md, const Real t_end)",
                "TaskStatus SourcePhotons(T *md, const Real t_start, const Real dt)",
                "TaskStatus UpdateDerivedTransportFields(MeshData<Real> *md, const Real dt)",
                "TaskStatus EvaluateRadiationEnergy(T *md)",
                "TaskStatus UpdateFluid(MeshData<Real> *md)",
                "TaskStatus DefragParticles(MeshBlock *)",
                "void PhotonReflectBC(std::shared_ptr<Swarm> &)"):
        assert sym in src, sym
    # every <jaybenne> input key of the reference, with its default
    for key in ("num_particles", "dt", "min_swarm_occupancy", "numin", "numax", "unique_rank_seeds", "seed",
                "max_transport_iterations", "use_ddmc", "tau_ddmc", "source_strategy", "do_emission", "do_feedback"):
        assert re.search(r'"jaybenne", "%s"' % key, src), key
    # no elided bodies
    assert "/* ..." not in src and "fill ..." not in src


def test_adapter_uses_only_declared_abi_entry_points():
    import sys
    sys.path.insert(0, ROOT)
    from jaybenne_b200 import capi
    used = set(re.findall(r"\b(jb_[a-z_0-9]+)\s*\(", open(ADAPTER).read()))
    assert used and used <= set(capi.declared_symbols()), used - set(capi.declared_symbols())
