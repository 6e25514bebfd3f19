function r = wrapTo360(a)
% SHIM, used only where the Mapping Toolbox / Octave `mapping` package is absent (run_reference_fixture adds this
% directory to the path only if `wrapTo360` does not resolve).  Semantics of the MathWorks function as documented:
% wrap to [0, 360], positive multiples of 360 map to 360, zero stays zero.
    r = mod(a, 360);
    r((r == 0) & (a > 0)) = 360;
end
