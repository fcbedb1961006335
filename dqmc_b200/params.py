"""`utility::parameters` of the reference (include/utility.h:50-276) for the Python host layer: the INI
dialect of examples/parameters.in with the same getters and failure behaviour.

Rules kept from the reference: `#` and `;` start a comment anywhere in a line; `[section]` headers
(default section "global"); `key = value`, both trimmed, one pair of matching quotes around the value
removed; numbers may contain `_` separators (10_000); booleans are true/1/yes/on and false/0/no/off in
any case; lists are comma separated with empty entries skipped.  A missing section/key or a value that
does not convert raises RuntimeError (std::runtime_error there); the getters with a default swallow
every failure and return the default, as the reference's catch-all does.
"""
from __future__ import annotations

_MISSING = object()


class Parameters:
    def __init__(self, filename: str | None = None, text: str | None = None):
        self.sections: dict[str, dict[str, str]] = {}
        if filename is not None:
            try:
                with open(filename) as f:
                    text = f.read()
            except OSError:
                raise RuntimeError(f"Failed to open parameter file: {filename}")
        section = "global"
        for line in (text or "").splitlines():
            section = self._parse_line(line, section)

    def _parse_line(self, line: str, section: str) -> str:
        for i, ch in enumerate(line):  # removeComment: everything after the first # or ;
            if ch in "#;":
                line = line[:i]
                break
        line = line.strip()
        if not line:
            return section
        if line[0] == "[" and line[-1] == "]":
            return line[1:-1].strip()
        if "=" in line:
            key, value = line.split("=", 1)
            key, value = key.strip(), value.strip()
            if value and ((value[0] == '"' and value[-1] == '"') or (value[0] == "'" and value[-1] == "'")):
                value = value[1:-1]  # a lone quote character becomes the empty string, as substr(1, len-2) does
            self.sections.setdefault(section, {})[key] = value
        return section

    # --- getters ---------------------------------------------------------------------------------
    def set(self, section: str, key: str, value) -> "Parameters":
        self.sections.setdefault(section, {})[key] = str(value)
        return self

    def getString(self, section: str, key: str, default=_MISSING) -> str:
        try:
            if section not in self.sections:
                raise RuntimeError(f"Section '{section}' not found")
            if key not in self.sections[section]:
                raise RuntimeError(f"Key '{key}' not found in section '{section}'")
            return self.sections[section][key]
        except RuntimeError:
            if default is _MISSING:
                raise
            return default

    @staticmethod
    def _stoi(s: str) -> int:
        """std::stoi: optional whitespace and sign, then the longest run of decimal digits."""
        t = s.lstrip()
        j = 1 if t[:1] in "+-" else 0
        k = j
        while k < len(t) and t[k].isdigit():
            k += 1
        if k == j:
            raise ValueError(s)
        return int(t[:k])

    @staticmethod
    def _stod(s: str) -> float:
        """std::stod: the longest prefix that parses as a floating-point number."""
        t = s.lstrip()
        for k in range(len(t), 0, -1):
            try:
                return float(t[:k])
            except ValueError:
                continue
        raise ValueError(s)

    def getInt(self, section: str, key: str, default=_MISSING) -> int:
        try:
            value = self.getString(section, key)
            try:
                return self._stoi(value.replace("_", ""))
            except ValueError:
                raise RuntimeError(f"Cannot convert '{value}' to integer for key '{key}'")
        except RuntimeError:
            if default is _MISSING:
                raise
            return default

    def getDouble(self, section: str, key: str, default=_MISSING) -> float:
        try:
            value = self.getString(section, key)
            try:
                return self._stod(value.replace("_", ""))
            except ValueError:
                raise RuntimeError(f"Cannot convert '{value}' to double for key '{key}'")
        except RuntimeError:
            if default is _MISSING:
                raise
            return default

    def getBool(self, section: str, key: str, default=_MISSING) -> bool:
        try:
            value = self.getString(section, key).lower()
            if value in ("true", "1", "yes", "on"):
                return True
            if value in ("false", "0", "no", "off"):
                return False
            raise RuntimeError(f"Cannot convert '{value}' to boolean for key '{key}'")
        except RuntimeError:
            if default is _MISSING:
                raise
            return default

    def getDoubleVector(self, section: str, key: str) -> list[float]:
        out = []
        for item in self.getString(section, key).split(","):
            item = item.strip()
            if not item:
                continue
            try:
                out.append(self._stod(item.replace("_", "")))
            except ValueError:
                raise RuntimeError(f"Cannot convert '{item}' to double in list for key '{key}'")
        return out

    def hasSection(self, section: str) -> bool:
        return section in self.sections

    def hasKey(self, section: str, key: str) -> bool:
        return key in self.sections.get(section, {})
