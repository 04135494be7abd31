#!/bin/sh
# Offline stand-in for every external tool the provisioning path shells out to
# (triton, ssh-keygen, terraform, ansible-playbook, curl, ssh, date, sleep).
#
# One script, installed under each tool's name in a scratch bin/ directory that
# is put first on PATH; it dispatches on basename($0).  Behaviour is scripted by
# plain files under $TK8S_STUB_STATE, and every invocation is appended to
# $TK8S_STUB_STATE/calls.log as "<tool>|<cwd relative to the workspace>|<argv>".
#
# The contracts mirrored here are the ones the reference relies on:
#   triton env|networks|packages      setup.sh:210,257,259,536,541
#   ssh-keygen -E md5 -lf / -l -f / -R setup.sh:217-224,506
#   terraform get|apply|destroy -force setup.sh:156-157,501 and the local-exec
#       IP capture of terraform/{master,host}/main.tf:29-31
#   ansible-playbook -i hosts clusterUp.yml   setup.sh:113
#   curl / ssh / date / sleep in the dashboard wait loop   setup.sh:59-85
tool=$(basename "$0")
S=${TK8S_STUB_STATE:?TK8S_STUB_STATE not set}
root=${TK8S_STUB_ROOT:-$PWD}
case "$PWD" in
    "$root") rel=. ;;
    "$root"/*) rel=./${PWD#"$root"/} ;;
    *) rel=$PWD ;;
esac
printf '%s|%s|%s\n' "$tool" "$rel" "$*" >> "$S/calls.log"
# What the tool inherits is part of the boundary: every line of ./config exported
# verbatim, literal double quotes and all (setup.sh:543-549), plus what bash itself
# exports (SHLVL, PWD and OLDPWD following each `cd`, `_` = path of the command).
# The provisioning tools get their whole environment recorded, the others a digest
# of it.  Left out: this stub's own plumbing, and PYTHON* (how the test harness
# makes an uninstalled Python package importable).
sbx=${S%/state}
envdump() { env | LC_ALL=C sort | LC_ALL=C sed -e '/^TK8S_STUB_/d' -e '/^PYTHON/d' -e "s|$sbx|<SBX>|g"; }
case "$tool" in
    terraform|ansible-playbook) envdump | sed 's/^/  env:/' >> "$S/calls.log" ;;
    *) sum=$(envdump | md5sum); printf '  env~:%.12s\n' "$sum" >> "$S/calls.log" ;;
esac

# modelled latency (milliseconds) of a tool that sits on a network round trip:
# a file latency_ms/<tool>.  /bin/sleep by absolute path: `sleep` on PATH is a stub.
if [ -f "$S/latency_ms/$tool" ]; then
    ms=$(cat "$S/latency_ms/$tool")
    /bin/sleep "$(printf '%d.%03d' $((ms / 1000)) $((ms % 1000)))"
fi

sub=${1:-}
# a tool that reads the stdin it inherits (terraform asking for approval, ssh
# without -n): a file eat/<tool>-<first arg> holding the number of lines it takes.
# The script's remaining answers are what such a tool consumes.
if [ -f "$S/eat/$tool-$sub" ]; then
    want=$(cat "$S/eat/$tool-$sub")
    got=0
    while [ "$got" -lt "$want" ] && IFS= read -r eaten; do
        printf '  stdin:%s\n' "$eaten" >> "$S/calls.log"
        got=$((got + 1))
    done
fi

# scripted failure: a file fail/<tool>-<first arg> holding the exit status
if [ -f "$S/fail/$tool-$sub" ]; then
    msg="$S/fail/$tool-$sub.stderr"
    [ -f "$msg" ] && cat "$msg" >&2
    exit "$(cat "$S/fail/$tool-$sub")"
fi

# next value of a per-tool sequence: prints file <dir>/<n> (n = 1,2,...) or <dir>/default
seq_next() {
    n=0
    [ -f "$S/$1.n" ] && n=$(cat "$S/$1.n")
    n=$((n + 1))
    echo "$n" > "$S/$1.n"
    if [ -f "$S/$1/$n" ]; then
        cat "$S/$1/$n"
    elif [ -f "$S/$1/default" ]; then
        cat "$S/$1/default"
    else
        return 1
    fi
}

case "$tool" in
triton)
    case "$sub" in
        env) cat "$S/triton_env.txt" ;;
        networks) cat "$S/networks.txt" ;;
        packages) cat "$S/packages.txt" ;;
        *) echo "triton stub: unsupported: $*" >&2; exit 2 ;;
    esac
    ;;
ssh-keygen)
    case "$sub" in
        -E)
            # modern form: ssh-keygen -E md5 -lf <path>
            [ -f "$S/keygen_no_E" ] && { echo "ssh-keygen: illegal option -- E" >&2; exit 1; }
            path=$4
            [ -f "$path" ] || { echo "ssh-keygen: $path: No such file or directory" >&2; exit 255; }
            f="$S/keygen/$(basename "$path")"
            [ -f "$f" ] || { echo "$path is not a public key file." >&2; exit 255; }
            cat "$f"
            ;;
        -l)
            # legacy form: ssh-keygen -l -f <path>
            path=$3
            [ -f "$path" ] || { echo "ssh-keygen: $path: No such file or directory" >&2; exit 255; }
            f="$S/keygen_legacy/$(basename "$path")"
            [ -f "$f" ] || f="$S/keygen_sha/$(basename "$path")"
            [ -f "$f" ] || { echo "$path is not a public key file." >&2; exit 255; }
            cat "$f"
            ;;
        -R)
            echo "# Host $2 found: line 1"
            echo "$HOME/.ssh/known_hosts updated." >&2
            ;;
        *) echo "ssh-keygen stub: unsupported: $*" >&2; exit 2 ;;
    esac
    ;;
terraform)
    case "$sub" in
        get)
            echo "Get: file://$rel/master"
            ;;
        apply)
            [ -f "$S/terraform_no_ips" ] && { echo "Error applying plan: provisioning limit"; exit 0; }
            master_ip=10.0.0.1
            [ -f "$S/master_ip" ] && master_ip=$(cat "$S/master_ip")
            i=0
            # one IP per module block, in file order, the way local-exec appends them
            LC_ALL=C grep -a 'source = ' rancher.tf | while read -r _ _ src; do
                case "$src" in
                    '"master"') echo "$master_ip" >> masters.ip ;;
                    '"host"') i=$((i + 1)); echo "10.0.1.$i" >> hosts.ip ;;
                esac
            done
            echo "Apply complete! Resources: $(LC_ALL=C grep -a -c 'source = ' rancher.tf) added, 0 changed, 0 destroyed."
            ;;
        destroy)
            echo "Destroy complete!"
            ;;
        *) echo "terraform stub: unsupported: $*" >&2; exit 2 ;;
    esac
    ;;
ansible-playbook)
    echo "PLAY RECAP (stub) $*"
    if [ -f "$S/env_id" ]; then
        mkdir -p tmp
        cat "$S/env_id" > tmp/kubernetes_environment.id
    fi
    ;;
curl)
    seq_next curl || echo "<title>Kubernetes Dashboard</title>"
    ;;
ssh)
    # ssh -o StrictHostKeyChecking=no root@<ip> docker ps | "docker stop -t0 <id>"
    target=$3
    ip=${target#root@}
    case "$4" in
        docker)
            [ -f "$S/ssh/$ip.ps" ] && cat "$S/ssh/$ip.ps"
            ;;
        *) : ;;
    esac
    if [ -f "$S/ssh/$ip.exit" ]; then exit "$(cat "$S/ssh/$ip.exit")"; fi
    ;;
date)
    seq_next date || echo 1500000000
    ;;
sleep)
    :
    ;;
*)
    echo "stubtool: no behaviour for $tool" >&2
    exit 127
    ;;
esac
exit 0
