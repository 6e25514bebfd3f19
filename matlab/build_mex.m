% build_mex.m -- compile the MEX gateway against the C-ABI library (run from this directory).
% Requires libekfslam_b200.so (make -C ../ekf_slam_b200/csrc) and a CUDA-capable B200; there is no CPU path.
here = fileparts(mfilename('fullpath'));
inc = fullfile(here, '..', 'include');
libdir = fullfile(here, '..', 'ekf_slam_b200');
if exist('OCTAVE_VERSION', 'builtin')
    mkoctfile('--mex', ['-I' inc], fullfile(here, 'ekfslam_mex.c'), ['-L' libdir], '-lekfslam_b200', ...
              ['-Wl,-rpath,' libdir], '-o', fullfile(here, 'ekfslam_mex'));
else
    mex(['-I' inc], fullfile(here, 'ekfslam_mex.c'), ['-L' libdir], '-lekfslam_b200', ...
        ['LDFLAGS=$LDFLAGS -Wl,-rpath,' libdir], '-outdir', here);
end
