function d = atan2d(y, x)
% SHIM for Octave versions without atan2d (added to the path only if `atan2d` does not resolve).
    d = atan2(y, x) * (180 / pi);
end
