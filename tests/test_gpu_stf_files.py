"""GPU, file level: the reference's own `stf` CLI against `stf_gpu` -- the SAME unmodified sources with the call sites of the hot path
routed to libvrgpu on the compiler's command line (oracle/Makefile, target stf_gpu) -- on one Gadget-1
snapshot, comparing what north_star says must stay unchanged: .fof.grp (src/io.cxx:252-289), .catalog_groups,
.catalog_particles (+ .unbound) (:377-930), .hierarchy and .properties.

Settings: examples/sample_dmcosmological_run.cfg of the reference (values restated below), with ASCII output (the compiled
reference has no HDF5) and the group array file switched on.  Cases: the 3DFOF halo catalogue alone (BASELINE configs[0]), 6DFOF
field halos, the substructure search (SearchSubSub: GetVelocityDensity, BuildNumInGroup / BuildPGList from the routed call sites)
without its unbinding step -- every file identical -- and the stock sample, whose unbinding (CheckUnboundGroups with group flags
and with ireorder = 0) runs into the reference's quirk Q10 and is compared to a stated tolerance.

Canonicalisation: group ids of equal-size groups may come in another order (quirk Q5: the reference's order among equal sizes is
its heap's), so groups are matched through their smallest particle id; everything else is compared as written."""
import os
import subprocess

import numpy as np
import pytest

import gadget1
import vrsynth

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STF = os.path.join(ROOT, "oracle", "_ref", "stf")
STF_GPU = os.path.join(ROOT, "oracle", "_ref", "stf_gpu")

# examples/sample_dmcosmological_run.cfg, line by line (units, search, unbinding, properties)
SAMPLE = dict(
    Input_includes_dm_particle=1, Input_includes_gas_particle=0, Input_includes_star_particle=0, Input_includes_bh_particle=0,
    Cosmological_input=1, Length_input_unit_conversion_to_output_unit=1.0, Velocity_input_unit_conversion_to_output_unit=1.0,
    Mass_input_unit_conversion_to_output_unit=1.0, Gravity=43.0211349, Hubble_unit=100.0, Length_unit_to_kpc=1000.0, Velocity_to_kms=1.0,
    Mass_to_solarmass=1.0e10, Comoving_units=0, Particle_search_type=2, Baryon_searchflag=0, Search_for_substructure=1, Singlehalo_search=0,
    Keep_FOF=0, Minimum_size=20, Minimum_halo_size=35, FoF_Field_search_type=5, Halo_3D_linking_length=0.20, Cell_fraction=0.01, Grid_type=1,
    Nsearch_velocity=32, Nsearch_physical=256, Local_velocity_density_approximate_calculation=2, FoF_search_type=1, Iterative_searchflag=1,
    Outlier_threshold=2.5, Substructure_physical_linking_length=0.10, Velocity_ratio=2.0, Velocity_opening_angle=0.10,
    Velocity_linking_length=0.20, Significance_level=1.0, Iterative_threshold_factor=1.0, Iterative_linking_length_factor=2.0,
    Iterative_Vratio_factor=1.0, Iterative_ThetaOp_factor=1.0, Halo_core_search=2, Use_adaptive_core_search=0,
    Use_phase_tensor_core_growth=2, Halo_core_ellx_fac=0.7, Halo_core_ellv_fac=2.0, Halo_core_ncellfac=0.005, Halo_core_num_loops=8,
    Halo_core_loop_ellx_fac=0.75, Halo_core_loop_ellv_fac=1.0, Halo_core_loop_elln_fac=1.2, Halo_core_phase_significance=2.0,
    Unbind_flag=1, Unbinding_type=0, Allowed_kinetic_potential_ratio=0.95, Min_bound_mass_frac=0.65, Bound_halos=0,
    Keep_background_potential=1, Frac_pot_ref=1.0, Min_npot_ref=20, Kinetic_reference_frame_type=0, h_val=1.0, Omega_m=0.3,
    Omega_Lambda=0.7, Critical_density=1.0, Virial_density=200, Omega_b=0.0, Inclusive_halo_masses=3, Iterate_cm_flag=0,
    Sort_by_binding_energy=1, Reference_frame_for_properties=2, Write_group_array_file=1, Separate_output_files=0, Binary_output=0,
    Snapshot_value=7, Verbose=0)

CASES = {
    "config1_3dfof_halos": dict(Search_for_substructure=0, Inclusive_halo_masses=0),
    "field_6dfof": dict(Search_for_substructure=0, Inclusive_halo_masses=0, FoF_Field_search_type=3, Halo_6D_linking_length_factor=1.0,
                        Halo_6D_vel_linking_length_factor=1.25),
    # the substructure search without its unbinding step: everything SearchSubSub routes to the library except CheckUnboundGroups
    "sample_with_substructure_no_unbinding": dict(Unbind_flag=0),
    # the stock sample.  Its unbinding passes run into quirk Q10 (unbind.cxx:643,700 reads `oldnumingroup` uninitialised; the
    # serial build carries it from group to group, the library implements the intended per-group value -- pinned in
    # tests/test_gpu_dropin.py): a handful of marginal members differ, so this case is compared to a stated tolerance
    "sample_with_substructure": dict(),
}
Q10_TOLERANCE = 2e-3   # fraction of the grouped particles whose object may differ in the stock sample (measured: 92 of 121572 = 7.6e-4)


def _run(binary, workdir, snap, cfg, out):
    log = os.path.join(workdir, os.path.basename(out) + ".log")
    with open(log, "w") as f:
        rc = subprocess.run([binary, "-i", snap, "-I", "1", "-o", out, "-C", cfg], stdout=f, stderr=subprocess.STDOUT, cwd=workdir,
                            timeout=900).returncode
    assert rc == 0, f"{os.path.basename(binary)} exited with {rc}:\n" + "".join(open(log).readlines()[-25:])


def _ints(path, skip=0):
    return np.loadtxt(path, dtype=np.int64, skiprows=skip, ndmin=1)


def _catalog(out):
    """Groups as {smallest member pid: (bound member pids in file order, unbound member pids)} + sizes in file order."""
    head = open(out + ".catalog_groups").read().split()
    ng = int(head[2])
    vals = np.array(head[4:], dtype=np.int64)
    sizes, off, uoff = vals[:ng], vals[ng:2 * ng], vals[2 * ng:3 * ng]
    pids = _ints(out + ".catalog_particles", 2)
    upids = _ints(out + ".catalog_particles.unbound", 2)
    end = np.append(off[1:], len(pids))
    uend = np.append(uoff[1:], len(upids))
    groups = {}
    for g in range(ng):
        b, u = pids[off[g]:end[g]], upids[uoff[g]:uend[g]]
        assert len(b) + len(u) == sizes[g]
        groups[int(min(b.min() if len(b) else 2 ** 62, u.min() if len(u) else 2 ** 62))] = (b, u, g)
    return sizes, groups


def _properties(out):
    lines = open(out + ".properties").read().splitlines()
    names = lines[2].split()
    rows = np.array([[float(x) for x in l.split()] for l in lines[3:]])
    return names, rows


@pytest.mark.parametrize("case", list(CASES))
def test_stf_gpu_writes_the_files_stf_writes(tmp_path, case):
    if not (os.path.exists(STF) and os.path.exists(STF_GPU)):
        pytest.skip("oracle/_ref/stf and stf_gpu are built where /root/reference exists (make -C oracle ref)")
    nside = 64
    d = vrsynth.make_box(nside ** 3, 5)
    snap = str(tmp_path / "snap")
    gadget1.write_snapshot(snap, d["pos"], d["vel"], d["pid"], vrsynth.BOX, d["mass"][0])
    cfg = str(tmp_path / "run.cfg")
    with open(cfg, "w") as f:
        for k, v in {**SAMPLE, **CASES[case]}.items():
            f.write(f"{k}={v}\n")
    ref_out, gpu_out = str(tmp_path / "ref"), str(tmp_path / "gpu")
    _run(STF, str(tmp_path), snap, cfg, ref_out)
    _run(STF_GPU, str(tmp_path), snap, cfg, gpu_out)

    # .fof.grp: one id per particle in input order
    fr, fg = _ints(ref_out + ".fof.grp"), _ints(gpu_out + ".fof.grp")
    assert fr[0] == fg[0] == nside ** 3
    cr, sr = vrsynth.canonical_groups(fr[1:], d["pid"])
    cg, sg = vrsynth.canonical_groups(fg[1:], d["pid"])
    assert len(sr) == len(sg)
    if case == "sample_with_substructure":
        # objects matched through their smallest member: the same objects but for Q10's marginal members
        ndiff = int(np.count_nonzero((fr[1:] > 0) != (fg[1:] > 0)))
        first_r = {int(d["pid"][fr[1:] == g].min()): g for g in np.unique(fr[1:]) if g > 0}
        first_g = {int(d["pid"][fg[1:] == g].min()): g for g in np.unique(fg[1:]) if g > 0}
        common = first_r.keys() & first_g.keys()
        moved = sum(int(np.count_nonzero((fr[1:] == first_r[k]) != (fg[1:] == first_g[k]))) for k in common)
        ngrouped = int(np.count_nonzero(fr[1:] > 0))
        print(f"{case}: {len(sr)} objects, {len(common)} matched by smallest member, {moved} member differences in those, "
              f"{ndiff} grouped/ungrouped differences, {ngrouped} grouped particles")
        assert len(common) >= 0.98 * len(sr)
        assert moved + ndiff <= Q10_TOLERANCE * ngrouped
        sizes_r, _ = _catalog(ref_out)
        sizes_g, _ = _catalog(gpu_out)
        assert len(sizes_r) == len(sizes_g) and np.abs(np.sort(sizes_r) - np.sort(sizes_g)).sum() <= Q10_TOLERANCE * ngrouped
        return
    assert np.array_equal(cr, cg)

    # .catalog_groups / .catalog_particles: the same sizes in the same order, the same members in the same (binding energy) order
    sizes_r, groups_r = _catalog(ref_out)
    sizes_g, groups_g = _catalog(gpu_out)
    assert np.array_equal(sizes_r, sizes_g)
    assert groups_r.keys() == groups_g.keys()
    moved = 0
    row_of = np.zeros(len(sizes_r), dtype=np.int64)          # reference row -> stf_gpu row
    for key, (b, u, g) in groups_r.items():
        b2, u2, g2 = groups_g[key]
        assert np.array_equal(b, b2) and np.array_equal(u, u2), f"group with smallest pid {key}"
        assert sizes_r[g] == sizes_g[g2]
        row_of[g] = g2
        moved += int(g != g2)
    print(f"{case}: {len(sizes_r)} objects, {moved} of them in another row (equal sizes, Q5)")

    # .properties: row by row through the same matching, every column but the ids that name rows
    names, pr = _properties(ref_out)
    names_g, pg = _properties(gpu_out)
    assert names == names_g and pr.shape == pg.shape
    idcols = [i for i, nm in enumerate(names) if nm.split("(")[0] in ("ID", "hostHaloID")]
    keep = [i for i in range(len(names)) if i not in idcols]
    a, b = pr[:, keep], pg[row_of][:, keep]
    exact = np.array_equal(a, b)
    assert np.allclose(a, b, rtol=1e-9, atol=0.0, equal_nan=True), names[keep[int(np.argmax(np.max(np.abs(a - b), axis=0)))]]
    print(f"{case}: {len(keep)} property columns {'identical as written' if exact else 'equal to 1e-9'}")
    if moved == 0:
        assert open(ref_out + ".hierarchy").read() == open(gpu_out + ".hierarchy").read()
