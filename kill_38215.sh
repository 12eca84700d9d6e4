ssh test@127.0.0.1 "kill -9 2761"
ssh test@127.0.0.1 "kill -9 2762"
kill -9 2109
