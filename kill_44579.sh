ssh test@127.0.0.1 "kill -9 1818"
ssh test@127.0.0.1 "kill -9 1819"
kill -9 1591
