function run_reference_fixture(ref_dir, fixtures_dir, out_dir)
% RUN_REFERENCE_FIXTURE  Reference-side witness for the oracle of this repository (SURVEY 8(c) deliverable 4).
%
%   run_reference_fixture('/path/to/SamShue/EKF_SLAM', 'tests/golden/ref_inputs', 'tests/golden/ref_outputs')
%
% Runs the UNMODIFIED reference classes (EKF_SLAM.m:26-151, EKF_SLAM_UC.m:27-152, Correspondence.m:28-88 from
% ref_dir) on the committed golden inputs, one sub-directory of fixtures_dir per case, with a stub landmark_list
% (RefFixtureStub.m) in place of ROS + RANSAC, and writes for every case
%     x_final.csv  P_final.csv  s_final.csv  x_checkpoints.csv (one row per checkpoint, NaN padded)  n_landmarks.csv (per step)
% at 17 significant digits.  `python tests/golden/compare_reference_fixture.py tests/golden/ref_outputs` (or
% `pytest tests/test_oracle_cpu.py -k witness`) then compares them with what oracle/dense_literal.py produced for the
% same inputs: with those files present the oracle is pinned to the reference itself, not only to its source text.
%
% Works in MATLAB (needs the Mapping Toolbox for wrapTo360, or the shim) and in GNU Octave >= 6 (classdef handle
% objects).  Neither exists in the image this repository was built in, so this script has not been executed there.
% Only the two shims in matlab/octave_shims (wrapTo360, atan2d) may be added, and only when the name does not resolve.
    here = fileparts(mfilename('fullpath'));
    addpath(here);                       % RefFixtureStub
    addpath(ref_dir);                    % the reference, untouched
    if isempty(which('wrapTo360')) || isempty(which('atan2d'))
        addpath(fullfile(here, 'octave_shims'), '-end');
        fprintf('note: using shims from matlab/octave_shims for missing toolbox functions\n');
    end
    which_ekf = which('EKF_SLAM');
    assert(~isempty(strfind(which_ekf, ref_dir)), 'EKF_SLAM resolves to %s, not to the reference checkout', which_ekf);
    cases = dir(fixtures_dir);
    for c = 1:numel(cases)
        if ~cases(c).isdir || cases(c).name(1) == '.', continue; end
        in = fullfile(fixtures_dir, cases(c).name);
        out = fullfile(out_dir, cases(c).name);
        if ~exist(out, 'dir'), mkdir(out); end
        cls = strtrim(fileread(fullfile(in, 'class.txt')));
        u = load_csv(fullfile(in, 'u.csv'), 2);
        rows = load_csv(fullfile(in, 'obs_rows.csv'), 3);
        counts = load_csv(fullfile(in, 'obs_counts.csv'), 1);
        tindex = load_csv(fullfile(in, 'table_index.csv'), 1);
        tloc = load_csv(fullfile(in, 'table_loc.csv'), 2);
        ntab = load_csv(fullfile(in, 'n_table.csv'), 1);
        cps = load_csv(fullfile(in, 'checkpoint_steps.csv'), 1);
        nmax = 3 + 2 * numel(tindex);
        h = feval(cls);                  % EKF_SLAM() or EKF_SLAM_UC(), constructors as upstream
        xs = nan(numel(cps), nmax);
        nl = zeros(numel(counts), 1);
        r0 = 0;
        for t = 1:numel(counts)
            k = counts(t);
            stub = RefFixtureStub(rows(r0 + 1:r0 + k, :), tindex(1:ntab(t)), tloc(1:ntab(t), :));
            r0 = r0 + k;
            h.predict(u(t, :));                         % SLAM.m:110
            h.measure([], u(t, :), stub);               % SLAM.m:116
            nl(t) = (numel(h.x) - 3) / 2;
            ci = find(cps == t);
            if ~isempty(ci), xs(ci, 1:numel(h.x)) = h.x(:)'; end
        end
        save_csv(fullfile(out, 'x_final.csv'), h.x(:)');
        save_csv(fullfile(out, 'P_final.csv'), h.P);
        save_csv(fullfile(out, 's_final.csv'), h.s(:)');
        save_csv(fullfile(out, 'x_checkpoints.csv'), xs);
        save_csv(fullfile(out, 'n_landmarks.csv'), nl(:)');
        fprintf('%s (%s): %d steps, %d landmarks -> %s\n', cases(c).name, cls, numel(counts), nl(end), out);
    end
end

function M = load_csv(path, ncols)
    fid = fopen(path, 'r');
    assert(fid > 0, 'cannot open %s', path);
    txt = fread(fid, inf, 'uint8=>char')';
    fclose(fid);
    v = sscanf(strrep(txt, ',', ' '), '%f');
    M = reshape(v, ncols, [])';
end

function save_csv(path, M)
    fid = fopen(path, 'w');
    for i = 1:size(M, 1)
        fprintf(fid, '%.17g', M(i, 1));
        fprintf(fid, ',%.17g', M(i, 2:end));
        fprintf(fid, '\n');
    end
    fclose(fid);
end
