ssh test@127.0.0.1 "kill -9 9063"
ssh test@127.0.0.1 "kill -9 9064"
ssh test@127.0.0.1 "kill -9 9065"
kill -9 8873
