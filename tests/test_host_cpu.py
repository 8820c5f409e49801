"""Host-side mirror of the reference interfaces (no GPU needed): Args defaults, file formats, ids, substep rule."""
import os

import numpy as np

import smc_physics_b200.host as host

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_args_defaults_match_args_jl():
    a = host.Args()     # args.jl:5-27
    assert (a.expt_id, a.model_id, a.target_id, a.algorithm) == ("BulletTest_Sphere", "Modelv5", "Sphere", "SMC")
    assert (a.num_particles, a.rejuvenation_moves, a.save_particles, a.predict, a.prediction_timesteps) == \
        (20, 1, True, False, 90)
    # Float32 literals promoted to Float64: 0.05f0, 0.075f0
    assert a.observation_noise == float(np.float32(0.05)) and a.transition_noise == float(np.float32(0.075))
    assert a.init_vel_noise == 0.0


def test_output_ids_and_directories(tmp_path):
    a = host.Args()
    assert host.generate_noise_id(a) == "Init00_Obs05_Tra075"          # the directory names under Analysis/
    assert host.generate_inference_param_id(host.Args(num_particles=50, rejuvenation_moves=5)) == "Par50_Rej5"
    host.DATA_ROOT = str(tmp_path)
    try:
        d = host.make_directories("Exp1/Modelv5/Sphere/Init00_Obs05_Tra075/SMC/Par20_Rej1/Data")
    finally:
        host.DATA_ROOT = ""
    assert d.endswith("Analysis/Exp1/Modelv5/Sphere/Init00_Obs05_Tra075/SMC/Par20_Rej1/Data/") and os.path.isdir(d)


def test_substep_rule():
    # `while dt <= sim.step_dur` (PhyBullet.jl:82-88): 5 substeps at 1/60, 9 at 1/30
    assert host.CudaSim(step_dur=1 / 60).substeps == 5 and host.CudaSim(step_dur=1 / 30).substeps == 9


def test_read_obs_file_realflow_fixture():
    f = os.path.join(GOLD, "Data/RealFlowData/Exp1/Cube_Ela5_Var3.csv")
    pos, ori, vel, obs, T = host.read_obs_file(f, "RealFlow", "SMC")
    assert np.allclose(pos, [0.0829755416851, 0.0179003832093, 0.603385401083])
    assert np.allclose(ori, [275.134194259, 131.160517617, 146.748292588])
    assert np.allclose(vel, [-9.22738634843, 9.57562023093, -0.271082018036])
    assert T == 63 and obs.shape == (63, 3)
    assert np.allclose(obs[0], pos, atol=1e-6)          # RealFlow: row 1 of _observed IS the initial position
    assert np.allclose(obs[1], [-0.222991079092026, 0.335414618253708, 0.589004874229431])


def test_filename_filters_and_order():
    names = ["Cube_Ela5_Var3.csv", "Cube_Ela5_Var3_observed.csv", "Cube_Ela0_Var10.csv", "Cube_Ela0_Var2.csv",
             ".DS_Store", "batch_1.csv", "Cube_Ela2_Var1_predicted.csv"]
    kept = host.filter_unwanted_filenames(names)
    assert kept == ["Cube_Ela5_Var3.csv", "Cube_Ela0_Var10.csv", "Cube_Ela0_Var2.csv"]
    assert sorted(kept, key=host.trial_order) == ["Cube_Ela0_Var2.csv", "Cube_Ela0_Var10.csv", "Cube_Ela5_Var3.csv"]


def test_write_to_csv_schema_and_truncation(tmp_path):
    # schema and 5-digit truncation of fileio.jl:255-278; compare with a shipped output of the reference,
    # Data/BulletData/Modelv1/Exp1/Sphere/Inferences/inferences_Ela7_Var12.csv: "1,0.47662,120.20946,1,-0.21844,..."
    traj = np.zeros((2, 3, 2), np.float32)
    traj[0, :, 0] = [-0.218449, 0.101021, 0.486915]
    traj[1, :, 0] = [1e-5 * 1.5, -2.0, 0.5]
    traj[:, :, 1] = 0.123456789
    p = host.Particles(np.array([0.476629, 1.0]), np.array([120.209469, -3.0000001]), traj)
    f = tmp_path / "out.csv"
    host.write_to_csv(p, str(f))
    lines = f.read_text().strip().split("\n")
    assert lines[0] == "particle,elasticity,weight,frame,x,y,z"
    assert lines[1] == "1,0.47662,120.20946,1,-0.21844,0.10102,0.48691"
    assert lines[2] == "1,0.47662,120.20946,2,1.0e-5,-2.0,0.5"          # Julia's float printout
    assert lines[3] == "2,1.0,-3.0,1,0.12345,0.12345,0.12345"
    assert len(lines) == 1 + 2 * 2
